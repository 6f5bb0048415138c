/* include/gsrast.h -- C ABI of libgsrast.so, the B200-native (sm_100a)
 * differentiable Gaussian rasterizer.
 *
 * This is the drop-in boundary for the one hot path of
 * stu214634/gaussian-transformer: submodules/diff-gaussian-rasterization
 * (abbreviated DGR).  Each entry point states the reference interface it
 * replaces (file:line under /root/reference).  Plain pointers and sizes only:
 * no torch / pybind / C++ types cross this boundary.
 *
 * Conventions
 *   - every pointer named in gsr_inputs / the gradient arguments is a DEVICE
 *     pointer (what tensor.data_ptr() returns); NULL means "absent", exactly
 *     as an empty tensor does in the reference (DGR/cuda_rasterizer/forward.cu:205,241);
 *   - all arrays are dense fp32 in the reference's layouts
 *     (DGR/rasterize_points.cu:94-111): means3D (P,3), sh (P,M,3),
 *     colors_precomp (P,3), opacities (P,1), scales (P,3), rotations (P,4),
 *     cov3D_precomp (P,6), viewmatrix / projmatrix 16 floats read column-major
 *     (DGR/cuda_rasterizer/auxiliary.h:58-77), campos 3, bg 3;
 *   - the library never allocates or frees device memory and keeps no global
 *     state: the caller owns outputs and the three scratch buffers (sized by
 *     gsr_scratch_bytes), which carry everything backward needs -- the same
 *     contract as the reference's geomBuffer / binningBuffer / imgBuffer
 *     (DGR/rasterize_points.cu:27-33,73-78);
 *   - `stream` is a cudaStream_t passed as void* (0 = legacy default stream);
 *     all work is enqueued on it; calls are re-entrant across threads/devices;
 *   - return value 0 = success; otherwise gsr_last_error() (thread-local)
 *     describes the failure.  The Python wrapper raises RuntimeError, the
 *     exception type the reference's callers catch
 *     (train_stacked_transformer.py:392).
 *
 * Forward is two-phase because the per-instance scratch can only be sized once
 * num_rendered is known -- the reference has the same host round trip
 * (DGR/cuda_rasterizer/rasterizer_impl.cu:277-285).
 */
#ifndef GSRAST_H_
#define GSRAST_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define GSR_API __attribute__((visibility("default")))
#else
#define GSR_API
#endif

#define GSR_OK 0
#define GSR_ERR_INVALID 1  /* bad argument (shape, NULL, alignment) */
#define GSR_ERR_CUDA 2     /* a CUDA call or kernel failed */
#define GSR_ERR_OVERFLOW 3 /* num_rendered does not fit the reference's int */
#define GSR_ERR_PREFILTERED 4 /* a Gaussian was culled although prefiltered=1 */

/* Per-call arguments: the positional argument list of the reference's
 * _C.rasterize_gaussians (DGR/rasterize_points.h:18-38) minus the tensors the
 * library writes.  P = means3D.size(0); M = sh.size(1) (0 when sh is absent,
 * DGR/rasterize_points.cu:83-87); D = active SH degree; W, H = image size. */
typedef struct gsr_inputs {
  int32_t P, D, M, W, H;
  float tanfovx, tanfovy, scale_modifier;
  int32_t prefiltered, debug;
  const float* bg;
  const float* means3D;
  const float* sh;
  const float* colors_precomp;
  const float* opacities;
  const float* scales;
  const float* rotations;
  const float* cov3D_precomp;
  const float* viewmatrix;
  const float* projmatrix;
  const float* campos;
  /* Optional split SH storage (scene/gaussian_model.py:108-111 keeps _features_dc (P,1,3) and _features_rest
   * (P,M-1,3) and concatenates them for every render): when sh_rest is non-NULL, `sh` is the DC array and sh_rest
   * the rest; M is still the total number of coefficients.  NULL = `sh` is the concatenated (P,M,3) array. */
  const float* sh_rest;
  /* Extension: non-zero = opacities / scales / rotations are the RAW parameters _opacity /
   * _scaling / _rotation of scene/gaussian_model.py:95-106; preprocess_fwd applies sigmoid / exp / normalize itself
   * (bit-identical to torch's kernels) and the backward returns the gradients of the RAW parameters (the chain rule
   * is folded into preprocess_bwd) -- no separate activation pass in either direction.  Requires scales + rotations. */
  int32_t raw_params;
} gsr_inputs;

/* Library version (major*10000 + minor*100 + patch) and the SM architecture the
 * kernels were compiled for (100 = sm_100a). */
GSR_API int gsr_version(void);
GSR_API int gsr_compiled_arch(void);

/* Thread-local description of the last non-zero return on this thread. */
GSR_API const char* gsr_last_error(void);

/* Scratch sizes in bytes.  Replaces the required<T>() size queries the
 * reference runs inside its resize callbacks
 * (DGR/cuda_rasterizer/rasterizer_impl.h:67-72, rasterizer_impl.cu:225-240,283).
 * geom depends on P; image on W,H; binning on R (= num_rendered, pass 0 when
 * not known yet); bwd on P (per-Gaussian accumulators of the backward pass).
 * Any out pointer may be NULL.  All buffers must be 256-byte aligned. */
GSR_API int gsr_scratch_bytes(int32_t P, int64_t R, int32_t W, int32_t H, size_t* geom,
                      size_t* binning, size_t* image, size_t* bwd);

/* Forward, phase 1: preprocess (EWA projection, cov3D/cov2D, SH -> RGB, radii,
 * tile rectangles) + per-tile instance counts + tile ranges.
 * Replaces FORWARD::preprocess and the InclusiveSum / D2H copy of
 * CudaRasterizer::Rasterizer::forward (DGR/cuda_rasterizer/rasterizer_impl.cu:198-281,
 * forward.cu:155-256).
 *   radii        int32[P]  out (0 for culled Gaussians; fully written)
 *   host_result  int64[2]  PINNED HOST memory, written asynchronously on
 *                          `stream`: [0] = num_rendered, [1] = error flags
 *                          (bit 0: prefiltered violation).  The caller must
 *                          synchronise `stream` before reading it.
 * P == 0 is legal: nothing is launched and host_result is zeroed on return. */
GSR_API int gsr_forward_pre(const gsr_inputs* in, void* geom, void* image, int32_t* radii,
                    int64_t* host_result, void* stream);

/* Forward, phase 2: instance scatter into per-tile buckets, per-tile depth sort
 * (order identical to the reference's stable 64-bit tile|depth radix sort),
 * alpha compositing.  Replaces duplicateWithKeys, cub::DeviceRadixSort,
 * identifyTileRanges and FORWARD::render
 * (DGR/cuda_rasterizer/rasterizer_impl.cu:283-331, forward.cu:261-374).
 *   R          num_rendered as read from host_result[0]
 *   out_color  float[3*H*W] out, fully written (C + T*bg). */
GSR_API int gsr_forward_post(const gsr_inputs* in, const void* geom, void* binning,
                     void* image, int64_t R, float* out_color, void* stream);

/* Both phases in one call, without GPU idle time across the num_rendered round trip: phase 2 is launched
 * speculatively against a binning buffer sized for capacity_R list entries (gsr_scratch_bytes(P, capacity_R, ...));
 * its kernels compare the device-side num_rendered with capacity_R and do nothing when it does not fit.  Returns
 * once host_result is valid (the GPU may still be rendering).  The caller checks host_result[0]:
 *   <= capacity_R : done, out_color / image / binning are (being) written on `stream`;
 *   >  capacity_R : allocate gsr_scratch_bytes(P, R, ...) and call gsr_forward_post as usual.
 * capacity_R = 0 (or binning = NULL) runs phase 1 only. */
GSR_API int gsr_forward(const gsr_inputs* in, void* geom, void* binning, int64_t capacity_R, void* image,
                        int32_t* radii, int64_t* host_result, float* out_color, void* stream);

/* gsr_forward without any host synchronisation, for CUDA-graph capture and small scenes where the num_rendered round
 * trip and launch latency dominate (extension; the reference always waits for num_rendered,
 * DGR/cuda_rasterizer/rasterizer_impl.cu:281).  `result` (int64[2] in device memory or in pinned host memory, which
 * the device addresses directly) receives {num_rendered, flags} by a store from a kernel on `stream`.  capacity_R > 0 and binning are required, P > 0, debug = 0.  When
 * num_rendered > capacity_R nothing is rendered and all tile ranges are emptied, so gsr_backward called with
 * R = capacity_R is a harmless no-op; the caller detects the overflow from `result` after its next synchronisation
 * and repeats the step with a larger buffer. */
GSR_API int gsr_forward_capacity(const gsr_inputs* in, void* geom, void* binning, int64_t capacity_R, void* image,
                                 int32_t* radii, int64_t* result, float* out_color, void* stream);

/* Backward.  Replaces CudaRasterizer::Rasterizer::backward
 * (DGR/cuda_rasterizer/rasterizer_impl.cu:340-434; BACKWARD::render
 * backward.cu:399-557, computeCov2DCUDA :144-274, preprocessCUDA :346-396) and
 * the nine zero-filled gradient tensors of the torch glue
 * (DGR/rasterize_points.cu:151-159): every non-NULL output is FULLY written
 * (zeros where the reference leaves its zero-initialised tensors untouched),
 * so the caller may pass uninitialised memory.
 *   dL_dpix      float[3*H*W]  upstream gradient of out_color
 *   dL_dmeans3D  float[P*3]    dL_dmeans2D float[P*3] (z = 0)
 *   dL_dsh       float[P*M*3]  (NULL allowed: sh absent, or the caller rebuilds it from exchanged factors, below)
 *   dL_dcolors   float[P*3]    dL_dopacity float[P]
 *   dL_dscales   float[P*3]    dL_drotations float[P*4]
 *   dL_dcov3D    float[P*6]    dL_dconic float[P*4] ((P,2,2): xx, xy, -, yy)
 *   bwd_scratch  gsr_scratch_bytes(...).bwd bytes. */
GSR_API int gsr_backward(const gsr_inputs* in, const int32_t* radii, const void* geom,
                 const void* binning, const void* image, int64_t R,
                 const float* dL_dpix, float* dL_dmeans3D, float* dL_dmeans2D,
                 float* dL_dsh, float* dL_dcolors, float* dL_dopacity,
                 float* dL_dscales, float* dL_drotations, float* dL_dcov3D,
                 float* dL_dconic, void* bwd_scratch, void* stream);

/* The same backward in its two stages, for callers that start work between them (the data-parallel exchange below
 * launches the all-gather of a view's SH factors while stage 2 runs):
 *   gsr_backward_blend     BACKWARD::render (backward.cu:399-557): per-pixel gradients -> per-Gaussian accumulators
 *                          in bwd_scratch (zeroed here);
 *   gsr_backward_geometry  computeCov2DCUDA + preprocessCUDA (backward.cu:144-396): accumulators -> the outputs of
 *                          gsr_backward, same meaning and NULL rules.
 * gsr_backward(...) == gsr_backward_blend(...) followed by gsr_backward_geometry(...) with the same arguments. */
GSR_API int gsr_backward_blend(const gsr_inputs* in, const void* geom, const void* binning, const void* image,
                               int64_t R, const float* dL_dpix, void* bwd_scratch, void* stream);
GSR_API int gsr_backward_geometry(const gsr_inputs* in, const int32_t* radii, const void* geom,
                                  float* dL_dmeans3D, float* dL_dmeans2D, float* dL_dsh, float* dL_dcolors,
                                  float* dL_dopacity, float* dL_dscales, float* dL_drotations, float* dL_dcov3D,
                                  float* dL_dconic, void* bwd_scratch, void* stream);

/* gsr_backward_geometry with split SH gradients (in->sh_rest non-NULL): dL_dsh receives the DC rows (P,1,3) and
 * dL_dsh_rest the rest (P,M-1,3) -- what autograd would hand back through the torch.cat of get_features. */
GSR_API int gsr_backward_geometry_split(const gsr_inputs* in, const int32_t* radii, const void* geom,
                                        float* dL_dmeans3D, float* dL_dmeans2D, float* dL_dsh, float* dL_dsh_rest,
                                        float* dL_dcolors, float* dL_dopacity, float* dL_dscales,
                                        float* dL_drotations, float* dL_dcov3D, float* dL_dconic,
                                        void* bwd_scratch, void* stream);

/* Data-parallel SH gradient exchange (no reference counterpart: the reference is single-GPU; this serves the
 * view-sharded path of SURVEY.md section 8e).  For one view, dL/dsh(i,k,c) = basis_k(dir_i) * f(i,c) with
 * f = dL/dcolor whose forward-clamped channels are zeroed (DGR/cuda_rasterizer/backward.cu:20-139), a rank-1
 * product per Gaussian.  Ranks all-gather f (12 B per Gaussian per view) instead of all-reducing 12*M B per
 * Gaussian, call gsr_backward with dL_dsh = NULL, and rebuild the sum over all views locally:
 *   gsr_sh_color_factors:      factors float[P*3] out, from the dL_dcolors of gsr_backward and the geom buffer of
 *                              the same forward;
 *   gsr_backward_sh_factors:   the same factors read from bwd_scratch right after gsr_backward_blend;
 *   gsr_sh_grad_from_factors:  dL_dsh float[P*M*3] (+)= sum_v basis(normalize(means3D - campos[v])) (x) factors[v];
 *                              campos float[V*3] and factors float[V*P*3] are DEVICE arrays; accumulate = 0
 *                              overwrites (every element written), 1 adds to the existing content. */
GSR_API int gsr_sh_color_factors(int32_t P, const void* geom, const float* dL_dcolors, float* factors,
                                 void* stream);
GSR_API int gsr_backward_sh_factors(int32_t P, const void* geom, const void* bwd_scratch, float* factors,
                                    void* stream);
GSR_API int gsr_sh_grad_from_factors(int32_t P, int32_t D, int32_t M, int32_t V, const float* means3D,
                                     const float* campos, const float* factors, float* dL_dsh,
                                     int32_t accumulate, void* stream);

/* Mean squared distance to the 3 nearest neighbours of every point: replaces SimpleKNN::knn / distCUDA2
 * (submodules/simple-knn/simple_knn.cu:185-221, spatial.cu:15-26), which scene/gaussian_model.py:134 calls to
 * initialise the Gaussian scales.  Exact 3-NN with the reference's distance arithmetic: results are bit-identical.
 *   points      float[P*3] device     mean_dist2  float[P] device out
 *   scratch     gsr_knn_scratch_bytes(P) bytes of device memory. */
GSR_API int gsr_knn_scratch_bytes(int32_t P, size_t* bytes);
GSR_API int gsr_knn_dist2(int32_t P, const float* points, float* mean_dist2, void* scratch, void* stream);

/* Fused training loss on the rendered image and its gradient, the dL/dpixel that gsr_backward consumes:
 *   L1 = mean |img - gt|                                   utils/loss_utils.py:17-18
 *   SSIM: 11x11 Gaussian window (sigma 1.5), zero padding, per channel, mean          :24-63
 *   loss = (1 - lambda) * L1 + lambda * (1 - SSIM)                                     train.py:91-92
 * replacing five depthwise convolutions, ~15 elementwise kernels and their autograd replay by two passes.
 *   img, gt   float[C*H*W] device (gt gets no gradient)
 *   out3      float[3] device out: L1, SSIM, loss
 *   dL_dimg   float[C*H*W] device out, d loss / d img (NULL: values only)
 *   scratch   gsr_loss_scratch_bytes(C, H, W) bytes. */
GSR_API int gsr_loss_scratch_bytes(int32_t C, int32_t H, int32_t W, size_t* bytes);
GSR_API int gsr_l1_ssim_loss(int32_t C, int32_t H, int32_t W, const float* img, const float* gt,
                             float lambda_dssim, float* out3, float* dL_dimg, void* scratch, void* stream);
/* The gradient pass on its own, for autograd: after gsr_l1_ssim_loss(..., dL_dimg = NULL, scratch, ...) on the same
 * img / gt / lambda, writes dL_dimg = (*upstream) * d loss / d img; `upstream` is a DEVICE float (NULL = 1) and
 * `scratch` the untouched scratch of that forward call. */
GSR_API int gsr_l1_ssim_backward(int32_t C, int32_t H, int32_t W, const float* img, const float* gt,
                                 float lambda_dssim, const float* upstream, const void* scratch, float* dL_dimg,
                                 void* stream);

/* The parameter activations in front of the rasterizer, one pass per direction over 32 B per Gaussian: replaces
 * get_opacity = sigmoid(_opacity), get_scaling = exp(_scaling), get_rotation = normalize(_rotation)
 * (scene/gaussian_model.py:43-51, 95-106) and their autograd replay.  Same expressions as torch's CUDA kernels
 * (1 / (1 + expf(-x)); expf; x / max(||x||_2, 1e-12)).
 *   opacity_raw float[P]  scaling_raw float[P*3]  rotation_raw float[P*4] (16-byte aligned), outputs likewise;
 *   backward: dL_d<activated> in, dL_d<raw> out. */
GSR_API int gsr_activate(int32_t P, const float* opacity_raw, const float* scaling_raw, const float* rotation_raw,
                         float* opacities, float* scales, float* rotations, void* stream);
GSR_API int gsr_activate_backward(int32_t P, const float* opacity_raw, const float* scaling_raw,
                                  const float* rotation_raw, const float* dL_dopacities, const float* dL_dscales,
                                  const float* dL_drotations, float* dL_dopacity_raw, float* dL_dscaling_raw,
                                  float* dL_drotation_raw, void* stream);

/* Per-iteration densification statistics for the Gaussians the view saw (radii > 0): replaces the boolean-mask
 * updates of train.py:113-116 and GaussianModel.add_densification_stats (scene/gaussian_model.py:405-407) --
 * max_radii2D = max(max_radii2D, radii), xyz_gradient_accum += ||dL_dmeans2D.xy||_2, denom += 1 -- without their
 * nonzero() host synchronisation.  radii int32[P] and dL_dmeans2D float[P*3] as returned by the rasterizer;
 * max_radii2D, xyz_gradient_accum, denom float[P], updated in place. */
GSR_API int gsr_densification_stats(int32_t P, const int32_t* radii, const float* dL_dmeans2D, float* max_radii2D,
                                    float* xyz_gradient_accum, float* denom, void* stream);

/* Fused Adam step over up to 8 parameter groups in ONE launch: replaces the torch.optim.Adam(l, lr=0.0, eps=1e-15)
 * step of scene/gaussian_model.py:149-167 / train.py:126-128 (betas (0.9, 0.999) by default, no weight decay, no
 * amsgrad; torch/optim/adam.py op order, bias corrections in double on the host).  Arrays are device float32 and
 * updated in place; `step` is the 1-based step count of that group's state AFTER this update; lr is read per call
 * (the reference reschedules the xyz group's lr every iteration, gaussian_model.py:169-175). */
typedef struct gsr_adam_group {
  float* param;
  const float* grad;
  float* exp_avg;
  float* exp_avg_sq;
  int64_t n;
  int64_t step;
  double lr;
} gsr_adam_group;
/* beta1, beta2, eps, lr are doubles: torch forms 1 - beta and lr / (1 - beta1^t) in double before rounding to float32 */
GSR_API int gsr_adam_step(int32_t n_groups, const gsr_adam_group* groups, double beta1, double beta2, double eps,
                          void* stream);

/* Row compaction behind densify / clone / split / prune.  Replaces the boolean-mask indexing and
 * torch.cat of scene/gaussian_model.py:273-327 (_prune_optimizer, prune_points, cat_tensors_to_optimizer): one
 * index list per mask, then ONE launch that moves the rows of every parameter, Adam moment and statistics tensor.
 * gsr_compact_indices: idx_out[k] = add + (index of the k-th non-zero mask byte), ascending = the order of
 * tensor[mask]; the number of indices goes to count_out (device or pinned host int64).
 * gsr_gather_rows: for every table, dst[j, :] = src[idx[j], :] for j < n_rows; idx[j] < 0 writes a row of zeros
 * (fresh Adam moments); a table may bring its own index list (parameters and their moments differ in the new rows).  Bit-identical to the torch statements (pure data movement). */
typedef struct gsr_gather_table {
  const float* src;
  float* dst;
  const int32_t* idx; /* NULL: the call's shared index list */
  int32_t row_floats;
  int32_t reserved;
} gsr_gather_table;
GSR_API int gsr_compact_scratch_bytes(int32_t P, size_t* bytes);
GSR_API int gsr_compact_indices(int32_t P, const uint8_t* mask, int32_t* idx_out, int32_t add, int64_t* count_out,
                                void* scratch, void* stream);
/* gsr_compact_indices with built-in scratch: allocates NOTHING (no cudaMalloc / cudaFree, no allocating library call),
 * for callers that cannot allocate (CUDA graph capture, tight serving loops).  No counterpart in the reference.
 * Same arguments and results as gsr_compact_indices minus `scratch`; the
 * block counts live in a static device array of 16,400 bytes that every GPU loading the library holds until exit.
 * P > GSR_COMPACT_NOALLOC_MAX_P is rejected with GSR_ERR_INVALID.  All calls on a device share the array (it keeps
 * the last call's contents; nothing in it needs to start at zero): calls that overlap on the GPU race -- issue them
 * on one stream or order them with events. */
#define GSR_COMPACT_NOALLOC_MAX_P (1 << 24)
GSR_API int gsr_compact_indices_noalloc(int32_t P, const uint8_t* mask, int32_t* idx_out, int32_t add,
                                        int64_t* count_out, void* stream);
GSR_API int gsr_gather_rows(int32_t n_tables, const gsr_gather_table* tables, int64_t n_rows, const int32_t* idx,
                            void* stream);

/* Data-parallel exchange over NVLink peer memory (gsr_dp.PeerExchange; no reference counterpart -- the reference
 * is single-GPU, the semantic is the batch sum of train_stacked_transformer.py:203-222 across ranks).
 * gsr_peer_allreduce: in-place SUM over `world` ranks of n_floats floats (a multiple of 4).  peer_ptrs[p] = address, in
 * THIS process, of rank p's buffer (a symmetric allocation; peer_ptrs[rank] is the local one); multicast_ptr != 0: the
 * allocation's multicast address, the NVSwitch then reduces (multimem.ld_reduce / multimem.st).  The caller orders
 * the ranks: every rank's buffer complete before any rank calls this, and nobody reads the result before every rank's
 * call has finished (a device-side barrier on either side).
 * gsr_sh_grad_from_peers: dL_dsh (P,16,3) (+)= sum over V views of basis(normalize(means3D - campos_v)) (x) f_v, with
 * view_ptrs[v] = address of view v's record on the rank that rendered it: P*3 factor floats (gsr_backward_sh_factors)
 * followed by the view's 3 camera-position floats. */
GSR_API int gsr_peer_allreduce(int32_t world, int32_t rank, const uint64_t* peer_ptrs, uint64_t multicast_ptr,
                               int64_t n_floats, void* stream);
/* gsr_peer_push: all-gather by push -- this rank's block of n_floats floats (src) is stored into slot `rank` (slots
 * are slot_floats apart) of the gather buffer of every OTHER rank, peer_dst_ptrs[p] = address in this process of rank
 * p's gather buffer (symmetric allocation).  At most max_ctas CTAs (0 = 64), so that it can run beside another kernel. */
GSR_API int gsr_peer_push(int32_t world, int32_t rank, const uint64_t* peer_dst_ptrs, const float* src,
                          int64_t n_floats, int64_t slot_floats, int32_t max_ctas, void* stream);
GSR_API int gsr_sh_grad_from_peers(int32_t P, int32_t D, int32_t M, int32_t V, const float* means3D,
                                   const uint64_t* view_ptrs, float* dL_dsh, int32_t accumulate, void* stream);

/* The depth sort of the forward on its own: stable ascending sort of n (uint32 key, uint32 value) pairs over all 32
 * key bits; the values of the result are the ORIGINAL INDICES of the keys (vals on entry is ignored), i.e. the
 * "value = Gaussian id" use of cub::DeviceRadixSort::SortPairs in DGR/cuda_rasterizer/rasterizer_impl.cu:300-308
 * restricted to what the depth-ordered binning needs.  Result in keys / vals; keys_tmp / vals_tmp are ping-pong
 * buffers of n words; scratch: gsr_sort_scratch_bytes(n).  variant: 1 = onesweep.cuh, 0 = depth_sort.cuh. */
GSR_API int gsr_sort_scratch_bytes(int32_t n, size_t* bytes);
GSR_API int gsr_sort_pairs(int32_t n, uint32_t* keys, uint32_t* keys_tmp, uint32_t* vals, uint32_t* vals_tmp,
                           void* scratch, int32_t variant, void* stream);

/* present[i] = (view-space z > 0.2).  Replaces Rasterizer::markVisible /
 * checkFrustum (DGR/cuda_rasterizer/rasterizer_impl.cu:54-66,141-153). */
GSR_API int gsr_mark_visible(int32_t P, const float* means3D, const float* viewmatrix,
                     const float* projmatrix, uint8_t* present, void* stream);

/* Optional per-stage device timing (used by bench.py for the roofline numbers; the reference has no counterpart
 * beyond the CUDA-event pair of train.py:44-45).  Process-wide (autograd runs backward on its own thread), mutex
 * protected, off by default.  While enabled, every call brackets each stage with CUDA events on the caller's stream (no synchronisation);
 * gsr_profile_read synchronises those events, returns per-stage total milliseconds and launch counts accumulated
 * since the previous read, and recycles the events. */
enum {
  GSR_STAGE_PREPROCESS_FWD = 0, /* preprocess_fwd_kernel */
  GSR_STAGE_DEPTH_SORT = 1,     /* depth sort of the P Gaussians (depth_sort.cuh) */
  GSR_STAGE_BIN_COUNT = 2,      /* bin_count_cta_kernel (or bin_pass_kernel<false>), column_scan_kernel, tile_scan_kernel */
  GSR_STAGE_BIN_SCATTER = 3,    /* bin_pass_kernel<true> or bin_scatter_cta_kernel */
  GSR_STAGE_RENDER_FWD = 4,     /* render_fwd_kernel */
  GSR_STAGE_RENDER_BWD = 5,     /* render_bwd2_kernel<false, PAIR> (render_bwd_kernel with option render_bwd = 1) */
  GSR_STAGE_PREPROCESS_BWD = 6, /* preprocess_bwd_staged_kernel / preprocess_bwd_kernel */
  GSR_STAGE_COUNT = 7
};
GSR_API int gsr_profile_enable(int on);
/* Number of kernels this library has launched in the process since the last reset (every launch site counts
 * itself); bench.py reports it as gpu_launches.  No reference counterpart. */
GSR_API int64_t gsr_launch_count(int reset);
/* Development / test switches (process-wide).  "bin_cta": -1 = pick the scatter kernel by list size (default),
 * 0 / 1 = force bin_pass_kernel<true> / bin_scatter_cta_kernel; "render_bwd": 2 = two pixels per thread (default),
 * 1 = one, 3 = two + test counters (accepted pairs, and pairs whose accept/reject differs between the forward's expf
 * and the backward's ex2.approx alpha, as three uint64 behind the accumulators of bwd_scratch); with 2, tile grids of at
 * most "pair_tiles" tiles (default 2048) run the instantiation that takes two list entries per step of its hit loop
 * (small scenes with uneven lists: config C1 150 -> 130 us), 4 / 5 = that instantiation forced / never.  "render_fwd": 2 / 1 likewise for the forward blend; "pre_fwd": 1 = one tile per CTA (default), 2 = the persistent
 * preprocess kernel with bulk-asynchronous (TMA) staging, measured slower; "sort": 1 = onesweep depth sort (default), 0 = the histogram / row-scan / scatter
 * sort.  Initial values come from the environment variables GSR_BIN_CTA / GSR_RENDER_FWD / GSR_RENDER_BWD / GSR_PAIR_TILES / GSR_PRE_FWD / GSR_SORT, read once at load. */
GSR_API int gsr_set_option(const char* name, int32_t value);
/* Self-test of the forward blend's packed expf (render_fwd.cuh expf2_exact) against expf for every float whose bit
 * pattern lies in [lo_bits, hi_bits] and for its negation; *mismatches_device (device uint64) receives the count. */
GSR_API int gsr_selftest_expf(uint32_t lo_bits, uint32_t hi_bits, uint64_t* mismatches_device, void* stream);
GSR_API int gsr_profile_read(float* ms /* [GSR_STAGE_COUNT] */, int32_t* counts /* [GSR_STAGE_COUNT] */);

/* Decoded views into the scratch buffers of a finished forward (device
 * pointers into geom / image / binning) -- the counterpart of
 * GeometryState / ImageState / BinningState::fromChunk
 * (DGR/cuda_rasterizer/rasterizer_impl.cu:155-194), used by the parity tests
 * for stage-by-stage comparison. */
typedef struct gsr_views {
  const float* xy_depth_radius; /* float4[P]: mean2D.x, .y, depth, radius (int bits) */
  const float* conic_opacity;   /* float4[P] */
  const float* rgb_flags;       /* float4[P]: r, g, b, (tiles_touched<<3 | clamped) bits */
  const float* final_T;         /* float[N] */
  const uint32_t* n_contrib;    /* uint32[N] */
  const uint32_t* ranges;       /* uint2[T] */
  const uint32_t* tile_count;   /* uint32[T] */
  const uint32_t* point_list;   /* uint32[R] sorted Gaussian ids */
} gsr_views;
GSR_API int gsr_get_views(int32_t P, int64_t R, int32_t W, int32_t H, const void* geom,
                  const void* binning, const void* image, gsr_views* out);

#ifdef __cplusplus
}
#endif
#endif /* GSRAST_H_ */
