// csrc/gsrast_api.cu -- the C ABI of libgsrast.so (see include/gsrast.h) and the
// host-side orchestration of the kernels.  Stateless: no globals besides a
// thread-local error string and the opt-in profiler; no device allocation.
#include <cuda_runtime.h>

#include <cstdio>
#include <cstring>
#include <atomic>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/gsrast.h"
#include "binning.cuh"
#include "common.cuh"
#include "depth_sort.cuh"
#include "onesweep.cuh"
#include <algorithm>
#include <cstdlib>

#include "knn.cuh"
#include "loss.cuh"
#include "adam.cuh"
#include "activations.cuh"
#include "preprocess_bwd.cuh"
#include "preprocess_fwd.cuh"
#include "render_bwd.cuh"
#include "render_fwd.cuh"
#include "sh_exchange.cuh"
#include "peer_exchange.cuh"
#include "compact.cuh"

using namespace gsr;

namespace {

thread_local std::string g_err;

int fail(int code, const std::string& msg) {
  g_err = msg;
  return code;
}

int cuda_fail(cudaError_t e, const char* where) {
  return fail(GSR_ERR_CUDA, std::string(where) + ": " + cudaGetErrorString(e));
}

#define GSR_CUDA(call, where)                         \
  do {                                                \
    cudaError_t e__ = (call);                         \
    if (e__ != cudaSuccess) return cuda_fail(e__, where); \
  } while (0)

// After a launch: always catch launch-configuration errors; with debug, also
// synchronise so that a faulting stage is named (the reference's CHECK_CUDA,
// DGR/cuda_rasterizer/auxiliary.h:166-173).
#define GSR_STAGE(name, debug, stream)                                  \
  do {                                                                  \
    cudaError_t e__ = cudaGetLastError();                               \
    if (e__ != cudaSuccess) return cuda_fail(e__, name);                \
    if (debug) {                                                        \
      e__ = cudaStreamSynchronize(stream);                              \
      if (e__ != cudaSuccess) return cuda_fail(e__, name);              \
    }                                                                   \
  } while (0)

// ---- optional per-stage device timing (gsr_profile_*): process-wide (autograd runs backward on its own
// thread), mutex-protected, off by default
struct StageSpan { int stage; cudaEvent_t a, b; };
struct Profiler {
  std::atomic<bool> on{false};
  std::mutex mu;
  std::vector<StageSpan> spans;
  std::vector<cudaEvent_t> pool;
  cudaEvent_t get() {
    if (!pool.empty()) { cudaEvent_t e = pool.back(); pool.pop_back(); return e; }
    cudaEvent_t e = nullptr;
    cudaEventCreate(&e);
    return e;
  }
};
Profiler g_prof;
struct ScopedStage {
  bool live;
  cudaStream_t st;
  StageSpan sp;
  ScopedStage(int stage, cudaStream_t s) : live(g_prof.on.load()), st(s) {
    if (!live) return;
    sp.stage = stage;
    {
      std::lock_guard<std::mutex> lk(g_prof.mu);
      sp.a = g_prof.get();
      sp.b = g_prof.get();
    }
    cudaEventRecord(sp.a, st);
  }
  ~ScopedStage() {
    if (!live) return;
    cudaEventRecord(sp.b, st);
    std::lock_guard<std::mutex> lk(g_prof.mu);
    g_prof.spans.push_back(sp);
  }
};

bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// Development / test switches, read from the environment ONCE per process (not per call).
int env_switch(const char* name, int dflt) {
  const char* v = getenv(name);
  return (v && v[0] >= '0' && v[0] <= '9') ? atoi(v) : dflt;
}
std::atomic<int> g_force_bin_cta{env_switch("GSR_BIN_CTA", -1)};  // -1 = by list size, 0 / 1 = forced
// blend kernels: 1 = one pixel per thread, 2 = two (default).  Backward: with 2, tile grids of at most g_pair_tiles tiles
// take the PAIR instantiation (two list entries per step of the hit loop; measured on config C1, 1450 tiles with lists
// of up to 650 entries: 150 -> 130 us, neutral within 3 % on evenly filled small grids, slower on large ones);
// 4 = PAIR forced, 5 = PAIR off, 3 = the counting test kernel
std::atomic<int> g_render_bwd{env_switch("GSR_RENDER_BWD", 2)};
std::atomic<int> g_render_fwd{env_switch("GSR_RENDER_FWD", 2)};
std::atomic<int> g_pair_tiles{env_switch("GSR_PAIR_TILES", 2048)};
std::atomic<int> g_pre_fwd{env_switch("GSR_PRE_FWD", 1)};         // 1 = one tile per CTA (default), 2 = persistent bulk-copy kernel
std::atomic<int> g_sort{env_switch("GSR_SORT", 1)};               // 1 = onesweep.cuh (default), 0 = depth_sort.cuh
std::atomic<long long> g_launches{0};                             // kernels launched by this library (gsr_launch_count)
#define GSR_LAUNCHED(n) g_launches.fetch_add((n), std::memory_order_relaxed)

// Per-device one-time kernel attributes (opt-in shared memory above 48 KB); the per-call path only reads a flag.
int device_setup() {
  static std::atomic<int> done[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return cuda_fail(cudaGetLastError(), "cudaGetDevice");
  if (dev >= 0 && dev < 64 && done[dev].load(std::memory_order_acquire)) return GSR_OK;
  GSR_CUDA(cudaFuncSetAttribute(bin_count_cta_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)BINC_SMEM_MAX), "bin_count smem attribute");
  GSR_CUDA(cudaFuncSetAttribute(preprocess_fwd_bulk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, PFB_SMEM),
           "preprocess_fwd_bulk smem attribute");
  if (dev >= 0 && dev < 64) done[dev].store(1, std::memory_order_release);
  return GSR_OK;
}

int check_inputs(const gsr_inputs* in) {
  if (!in) return fail(GSR_ERR_INVALID, "gsr_inputs is NULL");
  if (in->P < 0 || in->W <= 0 || in->H <= 0) return fail(GSR_ERR_INVALID, "bad P/W/H");
  if (in->P == 0) return GSR_OK;
  if (!in->means3D) return fail(GSR_ERR_INVALID, "means3D must have dimensions (num_points, 3)");
  if (!in->opacities || !in->viewmatrix || !in->projmatrix || !in->bg)
    return fail(GSR_ERR_INVALID, "opacities / viewmatrix / projmatrix / bg must not be NULL");
  if (!in->colors_precomp) {
    // DGR/cuda_rasterizer/rasterizer_impl.cu:242-245 has the same requirement for
    // NUM_CHANNELS != 3; with 3 channels SHs are needed instead.
    if (!in->sh || in->M <= 0) return fail(GSR_ERR_INVALID, "either sh (M > 0) or colors_precomp is required");
    if (in->D < 0 || in->D > 3 || (in->D + 1) * (in->D + 1) > in->M)
      return fail(GSR_ERR_INVALID, "sh degree does not fit the number of coefficients");
    if (!in->campos) return fail(GSR_ERR_INVALID, "campos must not be NULL");
    if (in->sh_rest) {
      if (in->M != 16 && in->M != 4)
        return fail(GSR_ERR_INVALID, "split SH storage (sh_rest) is built for M = 16 and M = 4; concatenate otherwise");
      if (!aligned16(in->sh) || !aligned16(in->sh_rest))
        return fail(GSR_ERR_INVALID, "split SH storage: sh and sh_rest must be 16-byte aligned");
    }
  }
  if (!in->cov3D_precomp) {
    if (!in->scales || !in->rotations)
      return fail(GSR_ERR_INVALID, "either scales+rotations or cov3D_precomp is required");
    if (!aligned16(in->rotations)) return fail(GSR_ERR_INVALID, "rotations must be 16-byte aligned");
  } else if (in->raw_params) {
    return fail(GSR_ERR_INVALID, "raw_params needs scales + rotations (the raw _scaling / _rotation), not cov3D_precomp");
  }
  return GSR_OK;
}

Cam make_cam(const gsr_inputs* in) {
  Cam c;
  c.view = in->viewmatrix;
  c.proj = in->projmatrix;
  c.campos = in->campos;
  c.tanfovx = in->tanfovx;
  c.tanfovy = in->tanfovy;
  // DGR/cuda_rasterizer/rasterizer_impl.cu:222-223
  c.focal_y = in->H / (2.0f * in->tanfovy);
  c.focal_x = in->W / (2.0f * in->tanfovx);
  c.scale_modifier = in->scale_modifier;
  c.W = in->W;
  c.H = in->H;
  c.gx = (in->W + TILE - 1) / TILE;
  c.gy = (in->H + TILE - 1) / TILE;
  c.raw_params = in->raw_params != 0;
  return c;
}

}  // namespace

extern "C" {

int gsr_version(void) { return 100; }
int gsr_compiled_arch(void) { return 100; }
const char* gsr_last_error(void) { return g_err.c_str(); }

int gsr_set_option(const char* name, int32_t value) {
  if (!name) return fail(GSR_ERR_INVALID, "gsr_set_option: NULL name");
  if (!strcmp(name, "bin_cta")) g_force_bin_cta.store(value < 0 ? -1 : (value != 0));
  else if (!strcmp(name, "sort")) g_sort.store(value != 0);
  else if (!strcmp(name, "render_fwd")) g_render_fwd.store(value == 1 ? 1 : 2);
  else if (!strcmp(name, "pair_tiles")) g_pair_tiles.store(value < 0 ? 0 : value);
  else if (!strcmp(name, "pre_fwd")) g_pre_fwd.store(value == 1 ? 1 : 2);
  else if (!strcmp(name, "render_bwd")) g_render_bwd.store(value == 1 || (value >= 3 && value <= 5) ? value : 2);
  else return fail(GSR_ERR_INVALID, std::string("gsr_set_option: unknown option ") + name);
  return GSR_OK;
}

int gsr_selftest_expf(uint32_t lo_bits, uint32_t hi_bits, uint64_t* mismatches_device, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (!mismatches_device || hi_bits < lo_bits) return fail(GSR_ERR_INVALID, "gsr_selftest_expf: bad arguments");
  GSR_CUDA(cudaMemsetAsync(mismatches_device, 0, 8, stream), "selftest memset");
  expf_selftest_kernel<<<148 * 8, 256, 0, stream>>>(lo_bits, hi_bits, (unsigned long long*)mismatches_device), GSR_LAUNCHED(1);
  GSR_STAGE("expf_selftest", 0, stream);
  return GSR_OK;
}

int64_t gsr_launch_count(int reset) {
  return reset ? g_launches.exchange(0) : g_launches.load();
}

int gsr_profile_enable(int on) {
  g_prof.on.store(on != 0);
  return GSR_OK;
}

int gsr_profile_read(float* ms, int32_t* counts) {
  if (!ms || !counts) return fail(GSR_ERR_INVALID, "gsr_profile_read: NULL output");
  for (int k = 0; k < GSR_STAGE_COUNT; k++) { ms[k] = 0.f; counts[k] = 0; }
  std::lock_guard<std::mutex> lk(g_prof.mu);
  for (const StageSpan& sp : g_prof.spans) {
    GSR_CUDA(cudaEventSynchronize(sp.b), "profile event sync");
    float t = 0.f;
    GSR_CUDA(cudaEventElapsedTime(&t, sp.a, sp.b), "profile elapsed");
    ms[sp.stage] += t;
    counts[sp.stage] += 1;
    g_prof.pool.push_back(sp.a);
    g_prof.pool.push_back(sp.b);
  }
  g_prof.spans.clear();
  return GSR_OK;
}

int gsr_scratch_bytes(int32_t P, int64_t R, int32_t W, int32_t H, size_t* geom, size_t* binning, size_t* image,
                      size_t* bwd) {
  if (P < 0 || R < 0 || W <= 0 || H <= 0) return fail(GSR_ERR_INVALID, "gsr_scratch_bytes: bad sizes");
  const size_t N = (size_t)W * H;
  const size_t T = (size_t)((W + TILE - 1) / TILE) * ((H + TILE - 1) / TILE);
  if (geom) *geom = geom_bytes((size_t)P);
  if (binning) *binning = bin_bytes((size_t)R);
  if (image) *image = image_bytes(N, T, (size_t)P);
  if (bwd) *bwd = bwd_bytes((size_t)P);
  return GSR_OK;
}

int gsr_get_views(int32_t P, int64_t R, int32_t W, int32_t H, const void* geom, const void* binning,
                  const void* image, gsr_views* out) {
  if (!out) return fail(GSR_ERR_INVALID, "gsr_get_views: out is NULL");
  memset(out, 0, sizeof(*out));
  const size_t N = (size_t)W * H;
  const size_t T = (size_t)((W + TILE - 1) / TILE) * ((H + TILE - 1) / TILE);
  if (geom) {
    GeomView g = geom_view(const_cast<void*>(geom), (size_t)P);
    out->xy_depth_radius = (const float*)g.g0;
    out->conic_opacity = (const float*)g.g1;
    out->rgb_flags = (const float*)g.g2;
  }
  if (image) {
    ImageView im = image_view(const_cast<void*>(image), N, T);
    out->final_T = im.final_T;
    out->n_contrib = im.n_contrib;
    out->ranges = (const uint32_t*)im.ranges;
    out->tile_count = im.tile_count;
  }
  if (binning) {
    BinView b = bin_view(const_cast<void*>(binning), (size_t)R);
    out->point_list = b.point_list;
  }
  return GSR_OK;
}

int gsr_mark_visible(int32_t P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                     uint8_t* present, void* stream_) {
  (void)projmatrix;
  cudaStream_t stream = (cudaStream_t)stream_;
  if (P < 0) return fail(GSR_ERR_INVALID, "gsr_mark_visible: P < 0");
  if (P == 0) return GSR_OK;
  if (!means3D || !viewmatrix || !present) return fail(GSR_ERR_INVALID, "gsr_mark_visible: NULL pointer");
  mark_visible_kernel<<<(P + 255) / 256, 256, 0, stream>>>(P, means3D, viewmatrix, present), GSR_LAUNCHED(1);
  GSR_STAGE("mark_visible", 0, stream);
  return GSR_OK;
}

// copy_result = false (gsr_forward_capacity): num_rendered and the flags stay in ctrl[]; the caller's next kernel
// delivers them
static int forward_pre_impl(const gsr_inputs* in, void* geom, void* image, int32_t* radii, int64_t* host_result,
                            void* stream_, bool copy_result) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (int rc = check_inputs(in)) return rc;
  if (!host_result) return fail(GSR_ERR_INVALID, "host_result is NULL");
  if (in->P == 0) {
    host_result[0] = 0;
    host_result[1] = 0;
    return GSR_OK;
  }
  if (!geom || !image || !radii) return fail(GSR_ERR_INVALID, "geom / image / radii must not be NULL");
  const Cam cam = make_cam(in);
  const size_t N = (size_t)in->W * in->H, T = (size_t)cam.gx * cam.gy;
  GeomView g = geom_view(geom, (size_t)in->P);
  ImageView im = image_view(image, N, T);
  const BinPlan plan = make_bin_plan(in->P, cam.gx, cam.gy);
  if (plan.n_bands == 0) return fail(GSR_ERR_INVALID, "image wider than 32768 pixels is not supported");
  GSR_CUDA(cudaMemsetAsync(im.ctrl, 0, 4 * CTRL_WORDS, stream), "memset ctrl");
  const int K = (in->D + 1) * (in->D + 1);
  const float* sh_rest = in->colors_precomp ? nullptr : in->sh_rest;
  const bool vec_sh = in->sh && !sh_rest && !in->colors_precomp && ((K * 3) % 4 == 0) && ((in->M * 12) % 16 == 0) &&
                      aligned16(in->sh);
  const int blocks = (in->P + 255) / 256;
  {
  ScopedStage prof_(GSR_STAGE_PREPROCESS_FWD, stream);
  const bool stage_in = aligned16(in->means3D) && (!in->scales || aligned16(in->scales));
  const bool staged_sh = in->sh && !sh_rest && !in->colors_precomp && (in->M == 16 || in->M == 4) && aligned16(in->sh);
#define GSR_PF_LAUNCH(VEC, STG)                                                                                    \
  preprocess_fwd_kernel<VEC, STG><<<blocks, 256, 0, stream>>>(in->P, in->D, in->M, in->means3D, in->scales,       \
                                                               in->rotations, in->opacities, in->sh,               \
                                                               VEC ? nullptr : sh_rest, in->colors_precomp,        \
                                                               in->cov3D_precomp, cam, in->prefiltered, radii, g,  \
                                                               im.ctrl)
#define GSR_PFS_LAUNCH(MT, STG)                                                                                     \
  preprocess_fwd_staged_kernel<MT, STG><<<(in->P + PF_THREADS - 1) / PF_THREADS, PF_THREADS, 0, stream>>>(         \
      in->P, in->D, in->means3D, in->scales, in->rotations, in->opacities, in->sh, in->cov3D_precomp, cam,         \
      in->prefiltered, radii, g, im.ctrl)
  const bool bulk = staged_sh && in->M == 16 && stage_in && in->scales && in->rotations && !in->cov3D_precomp &&
                    aligned16(in->opacities) && in->P >= PF_THREADS && g_pre_fwd.load(std::memory_order_relaxed) == 2;
  if (bulk) {
    // full tiles: persistent kernel with bulk-asynchronous staging; the last partial tile: the one-tile kernel on
    // pointers advanced to it
    if (int rc = device_setup()) return rc;
    const int n_tiles = in->P / PF_THREADS, tail = in->P - n_tiles * PF_THREADS;
    preprocess_fwd_bulk_kernel<<<std::min(n_tiles, 148 * 3), PF_THREADS, PFB_SMEM, stream>>>(
        n_tiles, in->D, in->means3D, in->scales, in->rotations, in->opacities, in->sh, cam, in->prefiltered, radii, g,
        im.ctrl);
    if (tail > 0) {
      const size_t off = (size_t)n_tiles * PF_THREADS;
      GeomView gt = g;
      gt.g0 += off; gt.g1 += off; gt.g2 += off; gt.keys[0] += off;
      preprocess_fwd_staged_kernel<16, true><<<1, PF_THREADS, 0, stream>>>(
          tail, in->D, in->means3D + 3 * off, in->scales + 3 * off, in->rotations + 4 * off, in->opacities + off,
          in->sh + 48 * off, nullptr, cam, in->prefiltered, radii + off, gt, im.ctrl);
      GSR_LAUNCHED(1);
    }
  } else if (staged_sh && in->M == 16) { if (stage_in) GSR_PFS_LAUNCH(16, true); else GSR_PFS_LAUNCH(16, false); }
  else if (staged_sh) { if (stage_in) GSR_PFS_LAUNCH(4, true); else GSR_PFS_LAUNCH(4, false); }
  else if (vec_sh && stage_in) GSR_PF_LAUNCH(true, true);
  else if (vec_sh) GSR_PF_LAUNCH(true, false);
  else if (stage_in) GSR_PF_LAUNCH(false, true);
  else GSR_PF_LAUNCH(false, false);
#undef GSR_PF_LAUNCH
#undef GSR_PFS_LAUNCH
  GSR_LAUNCHED(1);
  GSR_STAGE("preprocess_fwd", in->debug, stream);
  }
  // depth sort of the P Gaussians: 4 stable 8-bit passes, result in keys[0] / vals[0]
  {
  ScopedStage prof_(GSR_STAGE_DEPTH_SORT, stream);
  if (g_sort.load(std::memory_order_relaxed) == 1) {
    GSR_LAUNCHED(os_sort_launch(g.keys, g.vals, in->P, g.cta_hist, stream));
    GSR_STAGE("depth_sort", in->debug, stream);
  } else {
    const int n_cta = (int)sort_ctas((size_t)in->P);
    for (int pass = 0; pass < 4; pass++) {
      const int src = pass & 1, dst = src ^ 1, shift = 8 * pass;
      radix_hist_kernel<<<n_cta, RS_THREADS, 0, stream>>>(g.keys[src], in->P, shift, n_cta, g.cta_hist);
      uint32_t* digit_total = g.cta_hist + (size_t)256 * n_cta;
      radix_rowscan_kernel<<<256, 256, 0, stream>>>(g.cta_hist, n_cta, digit_total);
      radix_scatter_kernel<<<n_cta, RS_THREADS, 0, stream>>>(g.keys[src], pass == 0 ? nullptr : g.vals[src], in->P,
                                                            shift, n_cta, g.cta_hist, digit_total, g.keys[dst],
                                                            g.vals[dst]);
      GSR_LAUNCHED(3);
      GSR_STAGE("depth_sort", in->debug, stream);
    }
  }
  }
  {
  ScopedStage prof_(GSR_STAGE_BIN_COUNT, stream);
  if (bin_parallel(cam.gx, cam.gy, plan)) {
    // parallel binning, counting half: one CTA per chunk, all tiles in shared memory
    const size_t smem = binc_smem(cam.gx, cam.gy);
    if (int rc = device_setup()) return rc;
    bin_count_cta_kernel<<<dim3(plan.n_chunks, binc_bands(cam.gx, cam.gy)), BINC_THREADS, smem, stream>>>(
        in->P, g.vals[0], g.g0, cam.gx, cam.gy, plan.chunk, g.rects, im.matrix), GSR_LAUNCHED(1);
  } else {
    // order-preserving serial walk (tile grids too large for a shared-memory table), counting half
    const int tasks = plan.n_chunks * plan.n_bands;
    const int bin_smem = BIN_WARPS * plan.band_tiles * 4;  // <= 32 KB (BAND_TILES_MAX): no opt-in needed
    bin_pass_kernel<false><<<(tasks + BIN_WARPS - 1) / BIN_WARPS, BIN_WARPS * 32, bin_smem, stream>>>(
        in->P, g.vals[0], g.g0, cam.gx, cam.gy, plan, im.matrix, im.ranges, nullptr, nullptr, nullptr, 0u), GSR_LAUNCHED(1);
  }
  GSR_STAGE("bin_count", in->debug, stream);
  column_scan_kernel<<<((int)T + 31) / 32, 1024, 0, stream>>>(im.matrix, plan.n_chunks, (int)T, im.tile_count), GSR_LAUNCHED(1);
  GSR_STAGE("column_scan", in->debug, stream);
  tile_scan_kernel<<<1, 1024, 0, stream>>>((int)T, im.tile_count, im.ranges, im.ctrl), GSR_LAUNCHED(1);
  GSR_STAGE("tile_scan", in->debug, stream);
  }
  // ctrl[CTRL_R_LO..HI] -> host_result[0] (little endian int64), flags -> [1]
  static_assert(CTRL_R_HI == CTRL_R_LO + 1 && CTRL_FLAGS == CTRL_R_LO + 2, "ctrl words 4..7 = R, flags, 0");
  if (copy_result)
    GSR_CUDA(cudaMemcpyAsync(host_result, im.ctrl + CTRL_R_LO, 16, cudaMemcpyDeviceToHost, stream), "copy R, flags");
  return GSR_OK;
}

int gsr_forward_pre(const gsr_inputs* in, void* geom, void* image, int32_t* radii, int64_t* host_result,
                    void* stream_) {
  return forward_pre_impl(in, geom, image, radii, host_result, stream_, true);
}

}  // extern "C"

namespace {
// Phase 2 of the forward.  speculative: R is an upper bound (the capacity of `binning` in list entries) and the
// kernels check the device-side num_rendered against it themselves.
int forward_post_impl(const gsr_inputs* in, const void* geom, void* binning, void* image, int64_t R, bool speculative,
                      float* out_color, cudaStream_t stream) {
  if (int rc = check_inputs(in)) return rc;
  if (!out_color) return fail(GSR_ERR_INVALID, "out_color is NULL");
  if (R < 0) return fail(GSR_ERR_INVALID, "R < 0");
  if (R > 0x7fffffffLL)
    return fail(GSR_ERR_OVERFLOW, "num_rendered exceeds 2^31-1 (the reference's int would overflow)");
  if (in->P == 0) {
    // DGR/rasterize_points.cu:66,80-114: with no Gaussians the image stays 0 (not bg)
    GSR_CUDA(cudaMemsetAsync(out_color, 0, sizeof(float) * NCH * (size_t)in->W * in->H, stream), "memset out");
    return GSR_OK;
  }
  if (!geom || !image || (R > 0 && !binning)) return fail(GSR_ERR_INVALID, "scratch buffer is NULL");
  const Cam cam = make_cam(in);
  const size_t N = (size_t)in->W * in->H, T = (size_t)cam.gx * cam.gy;
  GeomView g = geom_view(const_cast<void*>(geom), (size_t)in->P);
  ImageView im = image_view(image, N, T);
  BinView b = bin_view(binning, (size_t)R);
  const uint32_t* guard = speculative ? im.ctrl : nullptr;
  const uint32_t cap = (uint32_t)R;
  if (R > 0) {
    ScopedStage prof_(GSR_STAGE_BIN_SCATTER, stream);
    const BinPlan plan = make_bin_plan(in->P, cam.gx, cam.gy);
    {
      // order-preserving walk; when the parallel count ran it left the clipped rectangles in depth order
      const int tasks = plan.n_chunks * plan.n_bands;
      const int bin_smem = BIN_WARPS * plan.band_tiles * 4;
      // lists that outgrow the L2 (or chunks with long serial walks) take the CTA-cooperative, band-major kernel
      const int force_cta = g_force_bin_cta.load(std::memory_order_relaxed);
      const bool long_lists = (size_t)R * 4 > ((size_t)64 << 20) || plan.chunk > 1024;
      const bool cta = plan.chunk >= 256 && (force_cta >= 0 ? force_cta == 1 : long_lists);
      if (cta) {
        // Every band CTA scans its whole chunk, so taller bands mean less scanning but a larger cursor table per
        // warp: ~960 tiles per band (30 KB per CTA) measured best at both 1080p (8 rows: 552 -> 422 us at 3 M
        // Gaussians) and 3840x2160 (4 rows; 8 rows 2.37 -> 2.59 ms).
        BinPlan cp = plan;
        cp.rows_per_band = std::max(1, std::min({960 / cam.gx, cam.gy, 1023}));
        cp.n_bands = (cam.gy + cp.rows_per_band - 1) / cp.rows_per_band;
        cp.band_tiles = cp.rows_per_band * cam.gx;
        const int cta_smem = BIN_WARPS * cp.band_tiles * 4;  // <= ~20 KB: no opt-in needed
        bin_scatter_cta_kernel<<<cp.n_chunks * cp.n_bands, BIN_WARPS * 32, cta_smem, stream>>>(
            in->P, g.vals[0], g.g0, cam.gx, cam.gy, cp, im.matrix, im.ranges, b.point_list,
            bin_parallel(cam.gx, cam.gy, plan) ? g.rects : nullptr, guard, cap), GSR_LAUNCHED(1);
      } else {
        bin_pass_kernel<true><<<(tasks + BIN_WARPS - 1) / BIN_WARPS, BIN_WARPS * 32, bin_smem, stream>>>(
            in->P, g.vals[0], g.g0, cam.gx, cam.gy, plan, im.matrix, im.ranges, b.point_list,
            bin_parallel(cam.gx, cam.gy, plan) ? g.rects : nullptr, guard, cap), GSR_LAUNCHED(1);
      }
      GSR_STAGE("bin_scatter", in->debug, stream);
    }
  }
  ScopedStage prof_(GSR_STAGE_RENDER_FWD, stream);
  dim3 grid(cam.gx, cam.gy);
  if (g_render_fwd.load(std::memory_order_relaxed) == 1)
    render_fwd_kernel<<<grid, TILE_PIX, 0, stream>>>(im.ranges, b.point_list, in->W, in->H, g.g0, g.g1, g.g2,
                                                     in->colors_precomp, in->bg, im.final_T, im.n_contrib, out_color,
                                                     guard, cap), GSR_LAUNCHED(1);
  else
    render_fwd2_kernel<<<grid, RF2_THREADS, 0, stream>>>(im.ranges, b.point_list, in->W, in->H, g.g0, g.g1, g.g2,
                                                         in->colors_precomp, in->bg, im.final_T, im.n_contrib,
                                                         out_color, guard, cap), GSR_LAUNCHED(1);
  GSR_STAGE("render_fwd", in->debug, stream);
  return GSR_OK;
}

// one event per (host thread, device) for gsr_forward's "R is on the host" wait; created once, never re-created
cudaEvent_t forward_event() {
  constexpr int MAX_DEV = 64;
  thread_local cudaEvent_t ev[MAX_DEV] = {};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= MAX_DEV) return nullptr;
  if (ev[dev] == nullptr && cudaEventCreateWithFlags(&ev[dev], cudaEventDisableTiming) != cudaSuccess) return nullptr;
  return ev[dev];
}
}  // namespace

extern "C" {

int gsr_forward_post(const gsr_inputs* in, const void* geom, void* binning, void* image, int64_t R,
                     float* out_color, void* stream_) {
  return forward_post_impl(in, geom, binning, image, R, false, out_color, (cudaStream_t)stream_);
}

int gsr_forward(const gsr_inputs* in, void* geom, void* binning, int64_t capacity_R, void* image, int32_t* radii,
                int64_t* host_result, float* out_color, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (capacity_R < 0) return fail(GSR_ERR_INVALID, "gsr_forward: capacity_R < 0");
  if (capacity_R > 0x7fffffffLL) capacity_R = 0x7fffffffLL;
  if (int rc = gsr_forward_pre(in, geom, image, radii, host_result, stream_)) return rc;
  if (in->P == 0) return forward_post_impl(in, geom, binning, image, 0, false, out_color, stream);
  cudaEvent_t ev = forward_event();
  if (!ev) return fail(GSR_ERR_CUDA, "gsr_forward: cudaEventCreate failed");
  GSR_CUDA(cudaEventRecord(ev, stream), "record R event");
  if (capacity_R > 0 && binning) {
    // phase 2 goes out BEFORE the host knows num_rendered: no GPU idle time across the round trip
    if (int rc = forward_post_impl(in, geom, binning, image, capacity_R, true, out_color, stream)) {
      cudaEventSynchronize(ev);  // the async copy into the caller's host_result must not outlive this call
      return rc;
    }
  }
  GSR_CUDA(cudaEventSynchronize(ev), "wait for num_rendered");
  return GSR_OK;
}

// The forward without any host synchronisation (CUDA-graph capturable, small-scene path): phase 1, then phase 2
// against `capacity_R` list entries.  `result` receives {num_rendered, flags} by a store from a kernel on `stream`
// (device memory, or pinned host memory, which the device addresses directly); the caller looks at it whenever it
// next synchronises.  On overflow
// (num_rendered > capacity_R) nothing is rendered and every tile range is emptied, so a backward launched blindly
// with R = capacity_R is a no-op rather than an out-of-bounds walk.
int gsr_forward_capacity(const gsr_inputs* in, void* geom, void* binning, int64_t capacity_R, void* image, int32_t* radii,
                         int64_t* result, float* out_color, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (!in) return fail(GSR_ERR_INVALID, "gsr_inputs is NULL");
  if (in->P <= 0) return fail(GSR_ERR_INVALID, "gsr_forward_capacity: P must be > 0");
  if (capacity_R <= 0 || capacity_R > 0x7fffffffLL || !binning)
    return fail(GSR_ERR_INVALID, "gsr_forward_capacity: needs a binning buffer of 0 < capacity_R <= 2^31-1 entries");
  if (in->debug) return fail(GSR_ERR_INVALID, "gsr_forward_capacity: debug synchronises; use gsr_forward");
  if (int rc = device_setup()) return rc;
  if (!result) return fail(GSR_ERR_INVALID, "gsr_forward_capacity: result is NULL");
  if (int rc = forward_pre_impl(in, geom, image, radii, result, stream_, false)) return rc;
  const Cam cam = make_cam(in);
  const size_t N = (size_t)in->W * in->H, T = (size_t)cam.gx * cam.gy;
  ImageView im = image_view(image, N, T);
  // delivers {num_rendered, flags} with a store from the device (a copy node between two kernels of a captured chain
  // costs ~15 us of idle GPU) and empties the ranges on overflow
  void_ranges_on_overflow_kernel<<<((int)T + 255) / 256, 256, 0, stream>>>((int)T, im.ranges, im.ctrl,
                                                                          (uint32_t)capacity_R, (long long*)result), GSR_LAUNCHED(1);
  GSR_STAGE("void_ranges_on_overflow", 0, stream);
  return forward_post_impl(in, geom, binning, image, capacity_R, true, out_color, stream);
}

int gsr_backward_blend(const gsr_inputs* in, const void* geom, const void* binning, const void* image, int64_t R,
                       const float* dL_dpix, void* bwd_scratch, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (int rc = check_inputs(in)) return rc;
  if (in->P == 0) return GSR_OK;
  if (!geom || !image || !dL_dpix || !bwd_scratch || (R > 0 && !binning))
    return fail(GSR_ERR_INVALID, "gsr_backward: NULL buffer");
  const Cam cam = make_cam(in);
  const size_t N = (size_t)in->W * in->H, T = (size_t)cam.gx * cam.gy;
  GeomView g = geom_view(const_cast<void*>(geom), (size_t)in->P);
  ImageView im = image_view(const_cast<void*>(image), N, T);
  BinView b = bin_view(const_cast<void*>(binning), (size_t)R);
  BwdView bw = bwd_view(bwd_scratch, (size_t)in->P);
  GSR_CUDA(cudaMemsetAsync(bwd_scratch, 0, bwd_bytes((size_t)in->P), stream), "memset bwd accumulators");
  if (R > 0) {
    ScopedStage prof_(GSR_STAGE_RENDER_BWD, stream);
    dim3 grid(cam.gx, cam.gy);
    const int variant = g_render_bwd.load(std::memory_order_relaxed);
    if (variant == 1)
      render_bwd_kernel<<<grid, TILE_PIX, 0, stream>>>(im.ranges, b.point_list, in->W, in->H, in->bg, g.g0, g.g1, g.g2,
                                                       in->colors_precomp, im.final_T, im.n_contrib, dL_dpix, bw), GSR_LAUNCHED(1);
    else if (variant == 3)
      render_bwd2_kernel<true><<<grid, RB2_THREADS, 0, stream>>>(im.ranges, b.point_list, in->W, in->H, in->bg, g.g0,
                                                                 g.g1, g.g2, in->colors_precomp, im.final_T,
                                                                 im.n_contrib, dL_dpix, bw, bw.counters), GSR_LAUNCHED(1);
    else if (variant == 4 || (variant == 2 && (int)T <= g_pair_tiles.load(std::memory_order_relaxed)))
      render_bwd2_kernel<false, true><<<grid, RB2_THREADS, 0, stream>>>(im.ranges, b.point_list, in->W, in->H, in->bg,
                                                                        g.g0, g.g1, g.g2, in->colors_precomp,
                                                                        im.final_T, im.n_contrib, dL_dpix, bw, nullptr),
          GSR_LAUNCHED(1);
    else
      render_bwd2_kernel<false><<<grid, RB2_THREADS, 0, stream>>>(im.ranges, b.point_list, in->W, in->H, in->bg, g.g0,
                                                                  g.g1, g.g2, in->colors_precomp, im.final_T,
                                                                  im.n_contrib, dL_dpix, bw, nullptr), GSR_LAUNCHED(1);
    GSR_STAGE("render_bwd", in->debug, stream);
  }
  return GSR_OK;
}

}  // extern "C"

namespace {
// Backward stage 2: accumulators -> gradients of the 3-D parameters (fused preprocess backward).
int backward_geometry_impl(const gsr_inputs* in, const int32_t* radii, const void* geom, float* dL_dmeans3D,
                           float* dL_dmeans2D, float* dL_dsh, float* dL_dsh_rest, float* dL_dcolors,
                           float* dL_dopacity, float* dL_dscales, float* dL_drotations, float* dL_dcov3D,
                           float* dL_dconic, void* bwd_scratch, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (int rc = check_inputs(in)) return rc;
  if (in->P == 0) return GSR_OK;
  if (!radii || !geom || !bwd_scratch) return fail(GSR_ERR_INVALID, "gsr_backward: NULL buffer");
  if (!dL_dmeans3D || !dL_dmeans2D || !dL_dcolors || !dL_dopacity || !dL_dscales || !dL_drotations ||
      !dL_dcov3D || !dL_dconic)
    return fail(GSR_ERR_INVALID, "gsr_backward: NULL gradient output");
  const Cam cam = make_cam(in);
  GeomView g = geom_view(const_cast<void*>(geom), (size_t)in->P);
  BwdView bw = bwd_view(bwd_scratch, (size_t)in->P);
  BwdOut out;
  out.dL_dmeans3D = dL_dmeans3D;
  out.dL_dmeans2D = dL_dmeans2D;
  out.dL_dsh = (in->M > 0) ? dL_dsh : nullptr;
  out.dL_dsh_rest = dL_dsh_rest;
  const bool split = in->sh && in->sh_rest && !in->colors_precomp;
  if (split && out.dL_dsh && !dL_dsh_rest)
    return fail(GSR_ERR_INVALID, "gsr_backward: split SH storage needs dL_dsh_rest (gsr_backward_geometry_split)");
  out.dL_dcolors = dL_dcolors;
  out.dL_dopacity = dL_dopacity;
  out.dL_dscales = dL_dscales;
  out.dL_drot = dL_drotations;
  out.dL_dcov3D = dL_dcov3D;
  out.dL_dconic = dL_dconic;
  ScopedStage prof_(GSR_STAGE_PREPROCESS_BWD, stream);
  // staged (coalesced) variant: SH rows of 16 or 4 coefficients and 16-byte aligned arrays; else the generic one
  const bool staged = in->sh && (in->M == 16 || in->M == 4) && aligned16(in->sh) &&
                      (!out.dL_dsh || (aligned16(dL_dsh) && (!split || aligned16(dL_dsh_rest)))) &&
                      aligned16(dL_dmeans3D) && aligned16(dL_dmeans2D) &&
                      aligned16(dL_dcolors) && aligned16(dL_dscales) && aligned16(dL_dcov3D) &&
                      aligned16(dL_dconic) && aligned16(dL_drotations);
  const int pb_blocks = (in->P + PB_THREADS - 1) / PB_THREADS;
  if (split && !staged) return fail(GSR_ERR_INVALID, "split SH storage: gradient outputs must be 16-byte aligned");
  if (split && in->M == 16) {
    preprocess_bwd_staged_kernel<16, true><<<pb_blocks, PB_THREADS, PB_THREADS * 45 * 4, stream>>>(
        in->P, in->D, in->means3D, in->scales, in->rotations, g.g1, in->sh, in->sh_rest, in->cov3D_precomp, cam,
        radii, g.g2, bw, out), GSR_LAUNCHED(1);
  } else if (split) {
    preprocess_bwd_staged_kernel<4, true><<<pb_blocks, PB_THREADS, PB_THREADS * 9 * 4 + 16, stream>>>(
        in->P, in->D, in->means3D, in->scales, in->rotations, g.g1, in->sh, in->sh_rest, in->cov3D_precomp, cam,
        radii, g.g2, bw, out), GSR_LAUNCHED(1);
  } else if (staged && in->M == 16) {
    const int smem = PB_THREADS * 13 * 16;
    preprocess_bwd_staged_kernel<16, false><<<pb_blocks, PB_THREADS, smem, stream>>>(
        in->P, in->D, in->means3D, in->scales, in->rotations, g.g1, in->sh, nullptr, in->cov3D_precomp, cam, radii,
        g.g2, bw, out), GSR_LAUNCHED(1);
  } else if (staged) {
    const int smem = PB_THREADS * 3 * 16;
    preprocess_bwd_staged_kernel<4, false><<<pb_blocks, PB_THREADS, smem, stream>>>(
        in->P, in->D, in->means3D, in->scales, in->rotations, g.g1, in->sh, nullptr, in->cov3D_precomp, cam, radii,
        g.g2, bw, out), GSR_LAUNCHED(1);
  } else {
    preprocess_bwd_kernel<<<pb_blocks, PB_THREADS, 0, stream>>>(in->P, in->D, in->M, in->means3D, in->scales,
                                                                in->rotations, g.g1, in->sh, in->cov3D_precomp,
                                                                cam, radii, g.g2, bw, out), GSR_LAUNCHED(1);
  }
  GSR_STAGE("preprocess_bwd", in->debug, stream);
  return GSR_OK;
}
}  // namespace

extern "C" {

int gsr_backward_geometry(const gsr_inputs* in, const int32_t* radii, const void* geom, float* dL_dmeans3D,
                          float* dL_dmeans2D, float* dL_dsh, float* dL_dcolors, float* dL_dopacity,
                          float* dL_dscales, float* dL_drotations, float* dL_dcov3D, float* dL_dconic,
                          void* bwd_scratch, void* stream_) {
  return backward_geometry_impl(in, radii, geom, dL_dmeans3D, dL_dmeans2D, dL_dsh, nullptr, dL_dcolors, dL_dopacity,
                                dL_dscales, dL_drotations, dL_dcov3D, dL_dconic, bwd_scratch, stream_);
}

int gsr_backward_geometry_split(const gsr_inputs* in, const int32_t* radii, const void* geom, float* dL_dmeans3D,
                                float* dL_dmeans2D, float* dL_dsh, float* dL_dsh_rest, float* dL_dcolors,
                                float* dL_dopacity, float* dL_dscales, float* dL_drotations, float* dL_dcov3D,
                                float* dL_dconic, void* bwd_scratch, void* stream_) {
  return backward_geometry_impl(in, radii, geom, dL_dmeans3D, dL_dmeans2D, dL_dsh, dL_dsh_rest, dL_dcolors,
                                dL_dopacity, dL_dscales, dL_drotations, dL_dcov3D, dL_dconic, bwd_scratch, stream_);
}

int gsr_backward(const gsr_inputs* in, const int32_t* radii, const void* geom, const void* binning,
                 const void* image, int64_t R, const float* dL_dpix, float* dL_dmeans3D, float* dL_dmeans2D,
                 float* dL_dsh, float* dL_dcolors, float* dL_dopacity, float* dL_dscales, float* dL_drotations,
                 float* dL_dcov3D, float* dL_dconic, void* bwd_scratch, void* stream_) {
  if (in && in->P > 0 && !radii) return fail(GSR_ERR_INVALID, "gsr_backward: NULL buffer");
  if (in && in->P > 0 &&
      (!dL_dmeans3D || !dL_dmeans2D || !dL_dcolors || !dL_dopacity || !dL_dscales || !dL_drotations ||
       !dL_dcov3D || !dL_dconic))
    return fail(GSR_ERR_INVALID, "gsr_backward: NULL gradient output");
  if (int rc = gsr_backward_blend(in, geom, binning, image, R, dL_dpix, bwd_scratch, stream_)) return rc;
  return gsr_backward_geometry(in, radii, geom, dL_dmeans3D, dL_dmeans2D, dL_dsh, dL_dcolors, dL_dopacity,
                               dL_dscales, dL_drotations, dL_dcov3D, dL_dconic, bwd_scratch, stream_);
}

// The view's SH factors straight from the blend accumulators (between the two backward stages).
int gsr_backward_sh_factors(int32_t P, const void* geom, const void* bwd_scratch, float* factors, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (P < 0) return fail(GSR_ERR_INVALID, "gsr_backward_sh_factors: P < 0");
  if (P == 0) return GSR_OK;
  if (!geom || !bwd_scratch || !factors) return fail(GSR_ERR_INVALID, "gsr_backward_sh_factors: NULL pointer");
  GeomView g = geom_view(const_cast<void*>(geom), (size_t)P);
  BwdView bw = bwd_view(const_cast<void*>(bwd_scratch), (size_t)P);
  sh_acc_factors_kernel<<<(P + 255) / 256, 256, 0, stream>>>(P, g.g2, bw.acc0, factors), GSR_LAUNCHED(1);
  GSR_STAGE("sh_acc_factors", 0, stream);
  return GSR_OK;
}

int gsr_sh_color_factors(int32_t P, const void* geom, const float* dL_dcolors, float* factors, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (P < 0) return fail(GSR_ERR_INVALID, "gsr_sh_color_factors: P < 0");
  if (P == 0) return GSR_OK;
  if (!geom || !dL_dcolors || !factors) return fail(GSR_ERR_INVALID, "gsr_sh_color_factors: NULL pointer");
  GeomView g = geom_view(const_cast<void*>(geom), (size_t)P);
  sh_color_factors_kernel<<<(P + 255) / 256, 256, 0, stream>>>(P, g.g2, dL_dcolors, factors), GSR_LAUNCHED(1);
  GSR_STAGE("sh_color_factors", 0, stream);
  return GSR_OK;
}

int gsr_sh_grad_from_factors(int32_t P, int32_t D, int32_t M, int32_t V, const float* means3D, const float* campos,
                             const float* factors, float* dL_dsh, int32_t accumulate, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (P < 0 || V < 0 || M <= 0 || D < 0 || D > 3 || (D + 1) * (D + 1) > M)
    return fail(GSR_ERR_INVALID, "gsr_sh_grad_from_factors: bad P / V / D / M");
  if (P == 0) return GSR_OK;
  if (!means3D || !dL_dsh || (V > 0 && (!campos || !factors)))
    return fail(GSR_ERR_INVALID, "gsr_sh_grad_from_factors: NULL pointer");
  const int blocks = (P + PB_THREADS - 1) / PB_THREADS;
  if (M == 16 && aligned16(dL_dsh))
    sh_grad_from_factors_kernel<true><<<blocks, PB_THREADS, PB_THREADS * 13 * 16, stream>>>(
        P, D, M, V, means3D, campos, factors, dL_dsh, accumulate), GSR_LAUNCHED(1);
  else
    sh_grad_from_factors_kernel<false><<<blocks, PB_THREADS, 0, stream>>>(P, D, M, V, means3D, campos, factors,
                                                                         dL_dsh, accumulate), GSR_LAUNCHED(1);
  GSR_STAGE("sh_grad_from_factors", 0, stream);
  return GSR_OK;
}

// ---- row compaction for densify / clone / split / prune
int gsr_compact_scratch_bytes(int32_t P, size_t* bytes) {
  if (P < 0 || !bytes) return fail(GSR_ERR_INVALID, "gsr_compact_scratch_bytes: bad arguments");
  *bytes = align_up(4 * ((size_t)P / CP_BLOCK + 2)) + ALIGN;
  return GSR_OK;
}

int gsr_compact_indices(int32_t P, const uint8_t* mask, int32_t* idx_out, int32_t add, int64_t* count_out,
                        void* scratch, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (P < 0 || !count_out) return fail(GSR_ERR_INVALID, "gsr_compact_indices: bad arguments");
  if (P == 0) {
    GSR_CUDA(cudaMemsetAsync(count_out, 0, 8, stream), "compact count memset");
    return GSR_OK;
  }
  if (!mask || !idx_out || !scratch) return fail(GSR_ERR_INVALID, "gsr_compact_indices: NULL pointer");
  const int nb = (P + CP_BLOCK - 1) / CP_BLOCK;
  uint32_t* counts = (uint32_t*)scratch;
  compact_count_kernel<<<nb, CP_THREADS, 0, stream>>>(P, mask, counts), GSR_LAUNCHED(1);
  compact_offsets_kernel<<<1, 1024, 0, stream>>>(nb, counts, (unsigned long long*)count_out), GSR_LAUNCHED(1);
  compact_write_kernel<<<nb, CP_THREADS, 0, stream>>>(P, mask, counts, idx_out, add), GSR_LAUNCHED(1);
  GSR_STAGE("compact_indices", 0, stream);
  return GSR_OK;
}

// ---- gsr_compact_indices without caller scratch and without any allocation (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: CUDA graph capture, serving loops).
// The block counts live in a file-scope device array that every GPU loading the library holds until exit
// (GSR_NOALLOC_SCRATCH_WORDS * 4 = 16,400 bytes).  All calls on one device share it: calls that overlap on the GPU
// (different streams) race.  Nothing in it has to start at zero -- compact_count_kernel writes every word that the two
// later kernels read -- so the previous call's contents are harmless.
constexpr int32_t GSR_NOALLOC_MAX_P = 1 << 24;                            // == GSR_COMPACT_NOALLOC_MAX_P (gsrast.h)
constexpr int GSR_NOALLOC_SCRATCH_WORDS = GSR_NOALLOC_MAX_P / CP_BLOCK + 4;
static_assert(GSR_NOALLOC_MAX_P == GSR_COMPACT_NOALLOC_MAX_P, "header and library disagree on the no-alloc maximum");
static __device__ __align__(256) uint32_t g_noalloc_block_counts[GSR_NOALLOC_SCRATCH_WORDS];

namespace {
// The array's name is only a host-side stand-in in host code: its device address is looked up once per device, with
// that device current, and cached.
int noalloc_scratch(uint32_t** out) {
  static std::mutex mu;
  static uint32_t* addr[64] = {};
  int dev = 0;
  GSR_CUDA(cudaGetDevice(&dev), "gsr_compact_indices_noalloc: cudaGetDevice");
  std::lock_guard<std::mutex> lk(mu);
  if (dev >= 0 && dev < 64 && addr[dev]) {
    *out = addr[dev];
    return GSR_OK;
  }
  uint32_t* p = nullptr;
  GSR_CUDA(cudaGetSymbolAddress((void**)&p, g_noalloc_block_counts), "gsr_compact_indices_noalloc: cudaGetSymbolAddress");
  if (!p) return fail(GSR_ERR_CUDA, "gsr_compact_indices_noalloc: cudaGetSymbolAddress returned NULL");
  if (dev >= 0 && dev < 64) addr[dev] = p;
  *out = p;
  return GSR_OK;
}
}  // namespace

int gsr_compact_indices_noalloc(int32_t P, const uint8_t* mask, int32_t* idx_out, int32_t add, int64_t* count_out,
                                void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (P < 0 || !count_out) return fail(GSR_ERR_INVALID, "gsr_compact_indices_noalloc: bad arguments");
  if (P > GSR_NOALLOC_MAX_P)
    return fail(GSR_ERR_INVALID, "gsr_compact_indices_noalloc: P = " + std::to_string(P) + " exceeds the built-in scratch (" +
                                     std::to_string(GSR_NOALLOC_MAX_P) + " mask entries); use gsr_compact_indices");
  if (P == 0) {
    GSR_CUDA(cudaMemsetAsync(count_out, 0, 8, stream), "compact count memset");
    return GSR_OK;
  }
  if (!mask || !idx_out) return fail(GSR_ERR_INVALID, "gsr_compact_indices_noalloc: NULL pointer");
  uint32_t* counts = nullptr;
  if (int rc = noalloc_scratch(&counts)) return rc;
  const int nb = (P + CP_BLOCK - 1) / CP_BLOCK;
  compact_count_kernel<<<nb, CP_THREADS, 0, stream>>>(P, mask, counts), GSR_LAUNCHED(1);
  compact_offsets_kernel<<<1, 1024, 0, stream>>>(nb, counts, (unsigned long long*)count_out), GSR_LAUNCHED(1);
  compact_write_kernel<<<nb, CP_THREADS, 0, stream>>>(P, mask, counts, idx_out, add), GSR_LAUNCHED(1);
  GSR_STAGE("compact_indices_noalloc", 0, stream);
  return GSR_OK;
}

int gsr_gather_rows(int32_t n_tables, const gsr_gather_table* tables, int64_t n_rows, const int32_t* idx,
                    void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (n_tables < 0 || n_rows < 0) return fail(GSR_ERR_INVALID, "gsr_gather_rows: bad arguments");
  if (n_tables == 0 || n_rows == 0) return GSR_OK;
  if (!tables) return fail(GSR_ERR_INVALID, "gsr_gather_rows: NULL pointer");
  for (int a0 = 0; a0 < n_tables; a0 += GATHER_MAX_TABLES) {
    GatherArgs a;
    memset(&a, 0, sizeof(a));
    long long blocks = 0;
    int k = 0;
    for (int i = a0; i < n_tables && i < a0 + GATHER_MAX_TABLES; i++) {
      const gsr_gather_table& s = tables[i];
      if (s.row_floats < 0) return fail(GSR_ERR_INVALID, "gsr_gather_rows: row_floats < 0");
      if (s.row_floats == 0) continue;
      if (!s.src || !s.dst) return fail(GSR_ERR_INVALID, "gsr_gather_rows: NULL table pointer");
      if (!s.idx && !idx) return fail(GSR_ERR_INVALID, "gsr_gather_rows: table without an index list");
      a.t[k].src = s.src; a.t[k].dst = s.dst; a.t[k].idx = s.idx; a.t[k].row_floats = s.row_floats; a.t[k].first_block = (int)blocks;
      blocks += (n_rows * s.row_floats + GATHER_ELEMS - 1) / GATHER_ELEMS;
      if (blocks > 0x7fffffffLL) return fail(GSR_ERR_INVALID, "gsr_gather_rows: too many elements for one launch");
      k++;
    }
    a.n_tables = k;
    if (k == 0) continue;
    gather_rows_kernel<<<(unsigned)blocks, GATHER_THREADS, 0, stream>>>(a, n_rows, idx), GSR_LAUNCHED(1);
  }
  GSR_STAGE("gather_rows", 0, stream);
  return GSR_OK;
}

// ---- data-parallel exchange over NVLink peer memory (gsr_dp.PeerExchange)
int gsr_peer_allreduce(int32_t world, int32_t rank, const uint64_t* peer_ptrs, uint64_t multicast_ptr, int64_t n_floats,
                       void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (world < 1 || world > PEER_MAX_RANKS || rank < 0 || rank >= world || n_floats < 0 || (n_floats & 3))
    return fail(GSR_ERR_INVALID, "gsr_peer_allreduce: bad world / rank / n_floats (a multiple of 4)");
  if (n_floats == 0 || world == 1) return GSR_OK;
  if (!peer_ptrs) return fail(GSR_ERR_INVALID, "gsr_peer_allreduce: NULL peer_ptrs");
  const size_t total4 = (size_t)n_floats / 4, n4 = (total4 + world - 1) / world;
  const int blocks = (int)std::min<size_t>((n4 + 255) / 256, (size_t)148 * 8);
  if (multicast_ptr) {
    const int mblocks = (int)std::min<size_t>((n4 + 1023) / 1024, (size_t)148 * 8);
    peer_allreduce_multimem_kernel<<<mblocks, 256, 0, stream>>>((float4*)multicast_ptr, rank, n4, total4), GSR_LAUNCHED(1);
  } else {
    PeerBufs b;
    memset(&b, 0, sizeof(b));
    for (int p = 0; p < world; p++) {
      if (!peer_ptrs[p] || (peer_ptrs[p] & 15u)) return fail(GSR_ERR_INVALID, "gsr_peer_allreduce: bad peer pointer");
      b.p[p] = (float4*)peer_ptrs[p];
    }
    switch (world) {
      case 2: peer_allreduce_kernel<2><<<blocks, 256, 0, stream>>>(b, rank, n4, total4), GSR_LAUNCHED(1); break;
      case 4: peer_allreduce_kernel<4><<<blocks, 256, 0, stream>>>(b, rank, n4, total4), GSR_LAUNCHED(1); break;
      case 8: peer_allreduce_kernel<8><<<blocks, 256, 0, stream>>>(b, rank, n4, total4), GSR_LAUNCHED(1); break;
      default: return fail(GSR_ERR_INVALID, "gsr_peer_allreduce: world must be 2, 4 or 8 without a multicast mapping");
    }
  }
  GSR_STAGE("peer_allreduce", 0, stream);
  return GSR_OK;
}

int gsr_peer_push(int32_t world, int32_t rank, const uint64_t* peer_dst_ptrs, const float* src, int64_t n_floats,
                  int64_t slot_floats, int32_t max_ctas, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (world < 1 || world > PEER_MAX_RANKS || rank < 0 || rank >= world || n_floats < 0 || (n_floats & 3) ||
      (slot_floats & 3) || slot_floats < n_floats)
    return fail(GSR_ERR_INVALID, "gsr_peer_push: bad world / rank / sizes (multiples of 4 floats, slot >= block)");
  if (n_floats == 0 || world == 1) return GSR_OK;
  if (!peer_dst_ptrs || !src || !aligned16(src)) return fail(GSR_ERR_INVALID, "gsr_peer_push: bad pointer");
  PeerBufs b;
  memset(&b, 0, sizeof(b));
  for (int p = 0; p < world; p++) {
    if (!peer_dst_ptrs[p] || (peer_dst_ptrs[p] & 15u)) return fail(GSR_ERR_INVALID, "gsr_peer_push: bad peer pointer");
    b.p[p] = (float4*)peer_dst_ptrs[p];
  }
  const size_t n4 = (size_t)n_floats / 4, slot4 = (size_t)slot_floats / 4;
  const int cap = max_ctas > 0 ? max_ctas : 64;
  const int blocks = (int)std::min<size_t>((n4 + 1023) / 1024, (size_t)cap);
  const float4* s4 = reinterpret_cast<const float4*>(src);
  switch (world) {
    case 2: peer_push_kernel<2><<<blocks, 256, 0, stream>>>(b, rank, s4, n4, slot4), GSR_LAUNCHED(1); break;
    case 4: peer_push_kernel<4><<<blocks, 256, 0, stream>>>(b, rank, s4, n4, slot4), GSR_LAUNCHED(1); break;
    case 8: peer_push_kernel<8><<<blocks, 256, 0, stream>>>(b, rank, s4, n4, slot4), GSR_LAUNCHED(1); break;
    default: return fail(GSR_ERR_INVALID, "gsr_peer_push: world must be 2, 4 or 8");
  }
  GSR_STAGE("peer_push", 0, stream);
  return GSR_OK;
}

int gsr_sh_grad_from_peers(int32_t P, int32_t D, int32_t M, int32_t V, const float* means3D,
                           const uint64_t* view_ptrs, float* dL_dsh, int32_t accumulate, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (P < 0 || V < 0 || V > PEER_MAX_VIEWS || M != 16 || D < 0 || D > 3)
    return fail(GSR_ERR_INVALID, "gsr_sh_grad_from_peers: bad P / V (<= 64) / D, or M != 16");
  if (P == 0) return GSR_OK;
  if (!means3D || !dL_dsh || (V > 0 && !view_ptrs)) return fail(GSR_ERR_INVALID, "gsr_sh_grad_from_peers: NULL pointer");
  if (!aligned16(dL_dsh)) return fail(GSR_ERR_INVALID, "gsr_sh_grad_from_peers: dL_dsh must be 16-byte aligned");
  PeerViews pv;
  memset(&pv, 0, sizeof(pv));
  for (int v = 0; v < V; v++) {
    if (!view_ptrs[v]) return fail(GSR_ERR_INVALID, "gsr_sh_grad_from_peers: NULL view pointer");
    pv.fac[v] = (const float*)view_ptrs[v];
  }
  const int blocks = (P + PB_THREADS - 1) / PB_THREADS;
  sh_grad_from_peers_kernel<8><<<blocks, PB_THREADS, PB_THREADS * 13 * 16, stream>>>(P, D, V, pv, means3D, dL_dsh,
                                                                                       accumulate), GSR_LAUNCHED(1);
  GSR_STAGE("sh_grad_from_peers", 0, stream);
  return GSR_OK;
}

// ---- the depth sort on its own (tests, tools/sort_bench.py)
int gsr_sort_scratch_bytes(int32_t n, size_t* bytes) {
  if (n < 0 || !bytes) return fail(GSR_ERR_INVALID, "gsr_sort_scratch_bytes: bad arguments");
  *bytes = align_up(4 * sort_scratch_words((size_t)n)) + ALIGN;
  return GSR_OK;
}

int gsr_sort_pairs(int32_t n, uint32_t* keys, uint32_t* keys_tmp, uint32_t* vals, uint32_t* vals_tmp, void* scratch,
                   int32_t variant, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (n < 0) return fail(GSR_ERR_INVALID, "gsr_sort_pairs: n < 0");
  if (n == 0) return GSR_OK;
  if (!keys || !keys_tmp || !vals || !vals_tmp || !scratch) return fail(GSR_ERR_INVALID, "gsr_sort_pairs: NULL pointer");
  uint32_t* k2[2] = {keys, keys_tmp};
  uint32_t* v2[2] = {vals, vals_tmp};
  if (variant == 1) {
    GSR_LAUNCHED(os_sort_launch(k2, v2, n, (uint32_t*)scratch, stream));
  } else {
    uint32_t* cta_hist = (uint32_t*)scratch;
    const int n_cta = (int)sort_ctas((size_t)n);
    for (int pass = 0; pass < 4; pass++) {
      const int src = pass & 1, dst = src ^ 1, shift = 8 * pass;
      radix_hist_kernel<<<n_cta, RS_THREADS, 0, stream>>>(k2[src], n, shift, n_cta, cta_hist);
      uint32_t* digit_total = cta_hist + (size_t)256 * n_cta;
      radix_rowscan_kernel<<<256, 256, 0, stream>>>(cta_hist, n_cta, digit_total);
      radix_scatter_kernel<<<n_cta, RS_THREADS, 0, stream>>>(k2[src], pass == 0 ? nullptr : v2[src], n, shift, n_cta,
                                                            cta_hist, digit_total, k2[dst], v2[dst]);
      GSR_LAUNCHED(3);
    }
  }
  GSR_STAGE("sort_pairs", 0, stream);
  return GSR_OK;
}

// ---- simple_knn.distCUDA2
namespace {
struct KnnView {
  uint32_t* bounds;
  uint32_t* keys[2];
  uint32_t* vals[2];
  uint32_t* cta_hist;
  float4* sorted;
  KnnBox* boxes;
  KnnBox* supers;
  size_t bytes;
};
KnnView knn_view(void* base, size_t P) {
  const size_t nb = (P + KNN_BOX - 1) / KNN_BOX, ns = (nb + KNN_SUPER - 1) / KNN_SUPER;
  char* p = (char*)base;
  KnnView v;
  v.bounds = (uint32_t*)p; p += ALIGN;
  for (int k = 0; k < 2; k++) { v.keys[k] = (uint32_t*)p; p += align_up(4 * P); }
  for (int k = 0; k < 2; k++) { v.vals[k] = (uint32_t*)p; p += align_up(4 * P); }
  v.cta_hist = (uint32_t*)p; p += align_up(4 * 256 * (sort_ctas(P) + 1));
  v.sorted = (float4*)p; p += align_up(16 * P);
  v.boxes = (KnnBox*)p; p += align_up(sizeof(KnnBox) * nb);
  v.supers = (KnnBox*)p; p += align_up(sizeof(KnnBox) * ns);
  v.bytes = (size_t)(p - (char*)base) + ALIGN;
  return v;
}
}  // namespace

int gsr_knn_scratch_bytes(int32_t P, size_t* bytes) {
  if (P < 0 || !bytes) return fail(GSR_ERR_INVALID, "gsr_knn_scratch_bytes: bad arguments");
  *bytes = knn_view(nullptr, (size_t)P).bytes;
  return GSR_OK;
}

int gsr_knn_dist2(int32_t P, const float* points, float* mean_dist2, void* scratch, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (P < 0) return fail(GSR_ERR_INVALID, "gsr_knn_dist2: P < 0");
  if (P == 0) return GSR_OK;
  if (!points || !mean_dist2 || !scratch) return fail(GSR_ERR_INVALID, "gsr_knn_dist2: NULL pointer");
  KnnView v = knn_view(scratch, (size_t)P);
  const int blocks = (P + 255) / 256;
  GSR_CUDA(cudaMemsetAsync(v.bounds, 0xff, 12, stream), "knn memset min");
  GSR_CUDA(cudaMemsetAsync(v.bounds + 3, 0x00, 12, stream), "knn memset max");
  knn_bounds_kernel<<<blocks < 592 ? blocks : 592, 256, 0, stream>>>(P, points, v.bounds), GSR_LAUNCHED(1);
  knn_morton_kernel<<<blocks, 256, 0, stream>>>(P, points, v.bounds, v.keys[0]), GSR_LAUNCHED(1);
  GSR_STAGE("knn_morton", 0, stream);
  const int n_cta = (int)sort_ctas((size_t)P);
  for (int pass = 0; pass < 4; pass++) {  // 30-bit codes: 4 stable 8-bit passes, result in keys[0] / vals[0]
    const int src = pass & 1, dst = src ^ 1, shift = 8 * pass;
    radix_hist_kernel<<<n_cta, RS_THREADS, 0, stream>>>(v.keys[src], P, shift, n_cta, v.cta_hist), GSR_LAUNCHED(1);
    uint32_t* digit_total = v.cta_hist + (size_t)256 * n_cta;
    radix_rowscan_kernel<<<256, 256, 0, stream>>>(v.cta_hist, n_cta, digit_total), GSR_LAUNCHED(1);
    radix_scatter_kernel<<<n_cta, RS_THREADS, 0, stream>>>(v.keys[src], pass == 0 ? nullptr : v.vals[src], P, shift,
                                                          n_cta, v.cta_hist, digit_total, v.keys[dst], v.vals[dst]), GSR_LAUNCHED(1);
  }
  GSR_STAGE("knn_sort", 0, stream);
  const int nb = (P + KNN_BOX - 1) / KNN_BOX, ns = (nb + KNN_SUPER - 1) / KNN_SUPER;
  knn_gather_kernel<<<blocks, 256, 0, stream>>>(P, points, v.vals[0], v.sorted), GSR_LAUNCHED(1);
  knn_boxes_kernel<<<nb, KNN_BOX, 0, stream>>>(P, v.sorted, v.boxes), GSR_LAUNCHED(1);
  knn_supers_kernel<<<ns, 32, 0, stream>>>(nb, v.boxes, v.supers), GSR_LAUNCHED(1);
  knn_search_kernel<<<nb, KNN_BOX, 0, stream>>>(P, v.sorted, nb, v.boxes, ns, v.supers, mean_dist2), GSR_LAUNCHED(1);
  GSR_STAGE("knn_search", 0, stream);
  return GSR_OK;
}

// ---- fused L1 + SSIM training loss
int gsr_loss_scratch_bytes(int32_t C, int32_t H, int32_t W, size_t* bytes) {
  if (C <= 0 || H <= 0 || W <= 0 || !bytes) return fail(GSR_ERR_INVALID, "gsr_loss_scratch_bytes: bad arguments");
  *bytes = align_up(LOSS_SLOTS * sizeof(LossAcc)) + align_up((size_t)3 * C * H * W * sizeof(float));
  return GSR_OK;
}

int gsr_l1_ssim_loss(int32_t C, int32_t H, int32_t W, const float* img, const float* gt, float lambda_dssim,
                     float* out3, float* dL_dimg, void* scratch, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (C <= 0 || H <= 0 || W <= 0) return fail(GSR_ERR_INVALID, "gsr_l1_ssim_loss: bad C / H / W");
  if (C > 65535) return fail(GSR_ERR_INVALID, "gsr_l1_ssim_loss: more than 65535 channels");
  if (!img || !gt || !out3 || !scratch) return fail(GSR_ERR_INVALID, "gsr_l1_ssim_loss: NULL pointer");
  LossAcc* acc = (LossAcc*)scratch;
  float* maps = (float*)((char*)scratch + align_up(LOSS_SLOTS * sizeof(LossAcc)));
  GSR_CUDA(cudaMemsetAsync(acc, 0, LOSS_SLOTS * sizeof(LossAcc), stream), "loss memset");
  const dim3 grid((W + LT - 1) / LT, (H + LT - 1) / LT, C);
  loss_pass1_kernel<<<grid, LOSS_THREADS, 0, stream>>>(H, W, img, gt, maps, acc), GSR_LAUNCHED(1);
  GSR_STAGE("loss_pass1", 0, stream);
  const double n = (double)C * H * W;
  loss_finish_kernel<<<1, 1, 0, stream>>>(acc, n, lambda_dssim, out3), GSR_LAUNCHED(1);
  if (dL_dimg) {
    loss_pass2_kernel<<<grid, LOSS_THREADS, 0, stream>>>(H, W, img, gt, maps, (float)((1.0 - lambda_dssim) / n),
                                                    (float)(lambda_dssim / n), nullptr, dL_dimg), GSR_LAUNCHED(1);
    GSR_STAGE("loss_pass2", 0, stream);
  }
  return GSR_OK;
}

int gsr_l1_ssim_backward(int32_t C, int32_t H, int32_t W, const float* img, const float* gt, float lambda_dssim,
                         const float* upstream, const void* scratch, float* dL_dimg, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (C <= 0 || H <= 0 || W <= 0 || C > 65535) return fail(GSR_ERR_INVALID, "gsr_l1_ssim_backward: bad C / H / W");
  if (!img || !gt || !scratch || !dL_dimg) return fail(GSR_ERR_INVALID, "gsr_l1_ssim_backward: NULL pointer");
  const float* maps = (const float*)((const char*)scratch + align_up(LOSS_SLOTS * sizeof(LossAcc)));
  const dim3 grid((W + LT - 1) / LT, (H + LT - 1) / LT, C);
  const double n = (double)C * H * W;
  loss_pass2_kernel<<<grid, LOSS_THREADS, 0, stream>>>(H, W, img, gt, maps, (float)((1.0 - lambda_dssim) / n),
                                                  (float)(lambda_dssim / n), upstream, dL_dimg), GSR_LAUNCHED(1);
  GSR_STAGE("loss_pass2", 0, stream);
  return GSR_OK;
}

// ---- parameter activations in front of the rasterizer
int gsr_activate(int32_t P, const float* opacity_raw, const float* scaling_raw, const float* rotation_raw,
                 float* opacities, float* scales, float* rotations, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (P < 0) return fail(GSR_ERR_INVALID, "gsr_activate: P < 0");
  if (P == 0) return GSR_OK;
  if (!opacity_raw || !scaling_raw || !rotation_raw || !opacities || !scales || !rotations)
    return fail(GSR_ERR_INVALID, "gsr_activate: NULL pointer");
  if (!aligned16(rotation_raw) || !aligned16(rotations))
    return fail(GSR_ERR_INVALID, "gsr_activate: rotation arrays must be 16-byte aligned");
  activate_kernel<<<(P + 255) / 256, 256, 0, stream>>>(P, opacity_raw, scaling_raw,
                                                       reinterpret_cast<const float4*>(rotation_raw), opacities, scales,
                                                       reinterpret_cast<float4*>(rotations)), GSR_LAUNCHED(1);
  GSR_STAGE("activate", 0, stream);
  return GSR_OK;
}

int gsr_activate_backward(int32_t P, const float* opacity_raw, const float* scaling_raw, const float* rotation_raw,
                          const float* dL_dopacities, const float* dL_dscales, const float* dL_drotations,
                          float* dL_dopacity_raw, float* dL_dscaling_raw, float* dL_drotation_raw, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (P < 0) return fail(GSR_ERR_INVALID, "gsr_activate_backward: P < 0");
  if (P == 0) return GSR_OK;
  if (!opacity_raw || !scaling_raw || !rotation_raw || !dL_dopacities || !dL_dscales || !dL_drotations ||
      !dL_dopacity_raw || !dL_dscaling_raw || !dL_drotation_raw)
    return fail(GSR_ERR_INVALID, "gsr_activate_backward: NULL pointer");
  if (!aligned16(rotation_raw) || !aligned16(dL_drotations) || !aligned16(dL_drotation_raw))
    return fail(GSR_ERR_INVALID, "gsr_activate_backward: rotation arrays must be 16-byte aligned");
  activate_bwd_kernel<<<(P + 255) / 256, 256, 0, stream>>>(
      P, opacity_raw, scaling_raw, reinterpret_cast<const float4*>(rotation_raw), dL_dopacities, dL_dscales,
      reinterpret_cast<const float4*>(dL_drotations), dL_dopacity_raw, dL_dscaling_raw,
      reinterpret_cast<float4*>(dL_drotation_raw)), GSR_LAUNCHED(1);
  GSR_STAGE("activate_bwd", 0, stream);
  return GSR_OK;
}

int gsr_densification_stats(int32_t P, const int32_t* radii, const float* dL_dmeans2D, float* max_radii2D,
                            float* xyz_gradient_accum, float* denom, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (P < 0) return fail(GSR_ERR_INVALID, "gsr_densification_stats: P < 0");
  if (P == 0) return GSR_OK;
  if (!radii || !dL_dmeans2D || !max_radii2D || !xyz_gradient_accum || !denom)
    return fail(GSR_ERR_INVALID, "gsr_densification_stats: NULL pointer");
  densify_stats_kernel<<<(P + 255) / 256, 256, 0, stream>>>(P, radii, dL_dmeans2D, max_radii2D, xyz_gradient_accum,
                                                            denom), GSR_LAUNCHED(1);
  GSR_STAGE("densify_stats", 0, stream);
  return GSR_OK;
}

// ---- fused multi-group Adam step
int gsr_adam_step(int32_t n_groups, const gsr_adam_group* groups, double beta1, double beta2, double eps,
                  void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (n_groups < 0 || n_groups > ADAM_MAX_GROUPS) return fail(GSR_ERR_INVALID, "gsr_adam_step: 0..8 groups per call");
  if (n_groups > 0 && !groups) return fail(GSR_ERR_INVALID, "gsr_adam_step: NULL groups");
  AdamArgs a;
  memset(&a, 0, sizeof(a));
  long long blocks = 0;
  int k = 0;
  for (int i = 0; i < n_groups; i++) {
    const gsr_adam_group& s = groups[i];
    if (s.n < 0 || s.step < 1) return fail(GSR_ERR_INVALID, "gsr_adam_step: n < 0 or step < 1");
    if (s.n == 0) continue;
    if (!s.param || !s.grad || !s.exp_avg || !s.exp_avg_sq) return fail(GSR_ERR_INVALID, "gsr_adam_step: NULL array");
    AdamGroup& g = a.grp[k++];
    g.p = s.param; g.g = s.grad; g.m = s.exp_avg; g.v = s.exp_avg_sq; g.n = s.n;
    // torch/optim/adam.py: bias_correction = 1 - beta ** step, step_size = lr / bias_correction1,
    // bias_correction2_sqrt = bias_correction2 ** 0.5 -- Python doubles, cast to float32 at the tensor op
    const double bc1 = 1.0 - pow(beta1, (double)s.step), bc2 = 1.0 - pow(beta2, (double)s.step);
    g.neg_step_size = (float)(-(s.lr / bc1));
    g.bc2_sqrt = (float)sqrt(bc2);
    g.first_block = (int)blocks;
    g.vec = aligned16(s.param) && aligned16(s.grad) && aligned16(s.exp_avg) && aligned16(s.exp_avg_sq);
    blocks += (s.n + ADAM_CHUNK - 1) / ADAM_CHUNK;
    if (blocks > 0x7fffffffLL) return fail(GSR_ERR_INVALID, "gsr_adam_step: too many elements for one launch");
  }
  if (blocks == 0) return GSR_OK;
  a.n_groups = k;
  a.w1 = (float)(1.0 - beta1);  // Python doubles in torch, rounded to float32 at the tensor op
  a.beta2 = (float)beta2;
  a.w2 = (float)(1.0 - beta2);
  a.eps = (float)eps;
  adam_step_kernel<<<(unsigned)blocks, ADAM_THREADS, 0, stream>>>(a), GSR_LAUNCHED(1);
  GSR_STAGE("adam_step", 0, stream);
  return GSR_OK;
}

}  // extern "C"
