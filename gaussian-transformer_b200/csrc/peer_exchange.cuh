// csrc/peer_exchange.cuh -- the data-parallel gradient exchange as OUR kernels over NVLink peer memory
// (gsr_dp.PeerExchange), instead of NCCL calls.
//
// The path shards over camera views: Gaussians are replicated, every rank renders its views, and the
// per-Gaussian gradients of the whole batch are needed on every rank.  Two kernels, both reading / writing the peers'
// buffers directly (symmetric allocations mapped into every rank's address space; NVSwitch gives every GPU full
// bandwidth to every peer, so there is no ring or tree to follow):
//
//   peer_allreduce_kernel   the 44 B/Gaussian small gradients (means3D, opacity, scales, rotations): two-shot
//                           all-reduce in ONE kernel -- rank r sums slice r of all ranks' buffers (one 128-bit load
//                           per peer in flight per thread) and stores the sum into slice r of every rank's buffer.
//                           Where the allocation has a multicast mapping, the switch does both halves
//                           (multimem.ld_reduce / multimem.st): 1/N of the link traffic.
//   sh_grad_from_peers      dL/dsh(i) = sum over the batch's views v of basis(dir_v(i)) (x) f_v(i) (sh_exchange.cuh)
//                           from a TABLE of per-view record addresses: own records in place, the peers' either in
//                           place over NVLink (gather fused into the rebuild) or -- what gsr_dp.PeerExchange does,
//                           measured faster at 8 GPUs -- from a staging copy that the copy engines filled while
//                           preprocess_bwd ran (the rebuild is HBM bound and then overlaps the link-bound all-reduce).
//
// Ordering between ranks (a rank may only read what a peer has finished writing, and may only overwrite what every
// peer has finished reading) is the caller's: gsr_dp.PeerExchange brackets the kernels with device-side barriers on
// the symmetric allocation's signal pads.  No reference counterpart (the reference is single-GPU).
#pragma once
#include "common.cuh"
#include "preprocess_fwd.cuh"
#include "preprocess_bwd.cuh"

namespace gsr {

constexpr int PEER_MAX_RANKS = 16;
constexpr int PEER_MAX_VIEWS = 64;

struct PeerBufs {
  float4* p[PEER_MAX_RANKS];
};
struct PeerViews {
  const float* fac[PEER_MAX_VIEWS];  // (P,3) factors of view v, then its 3 camera-position floats at fac[v] + 3 P
};

// n4 float4 per rank slice; rank r owns [r * n4, min((r + 1) * n4, total4)).
template <int WORLD>
__global__ void __launch_bounds__(256) peer_allreduce_kernel(PeerBufs bufs, int rank, size_t n4, size_t total4) {
  const size_t lo = (size_t)rank * n4;
  const size_t hi = lo + n4 < total4 ? lo + n4 : total4;
  for (size_t i = lo + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < hi; i += (size_t)gridDim.x * blockDim.x) {
    float4 v[WORLD];
#pragma unroll
    for (int p = 0; p < WORLD; p++) v[p] = bufs.p[p][i];  // independent loads: all peers in flight at once
    float4 s = v[0];
#pragma unroll
    for (int p = 1; p < WORLD; p++) { s.x += v[p].x; s.y += v[p].y; s.z += v[p].z; s.w += v[p].w; }
#pragma unroll
    for (int p = 0; p < WORLD; p++) bufs.p[p][i] = s;  // every rank receives the same bits
  }
}

// Same slices, through the allocation's multicast address: the NVSwitch reduces the ranks' copies on the load and
// replicates the store.
__global__ void __launch_bounds__(256) peer_allreduce_multimem_kernel(float4* mc, int rank, size_t n4, size_t total4) {
  const size_t lo = (size_t)rank * n4;
  const size_t hi = lo + n4 < total4 ? lo + n4 : total4;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  constexpr int U = 4;  // requests in flight per thread (a round trip through the switch is several microseconds)
  for (size_t i0 = lo + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i0 < hi; i0 += U * stride) {
    float4 s[U];
#pragma unroll
    for (int u = 0; u < U; u++) {
      const size_t i = i0 + u * stride;
      if (i < hi)
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                     : "=f"(s[u].x), "=f"(s[u].y), "=f"(s[u].z), "=f"(s[u].w)
                     : "l"(mc + i)
                     : "memory");
    }
#pragma unroll
    for (int u = 0; u < U; u++) {
      const size_t i = i0 + u * stride;
      if (i < hi)
        asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc + i), "f"(s[u].x),
                     "f"(s[u].y), "f"(s[u].z), "f"(s[u].w)
                     : "memory");
    }
  }
}

// All-gather by PUSH: this rank's block of n4 float4 is stored into slot `rank` of every peer's gather buffer
// (dst.p[p] = base of peer p's buffer; slots are slot4 float4 apart).  Stores over NVLink are posted, so a few CTAs
// saturate the links and the kernel can run beside preprocess_bwd; U loads in flight per thread.
template <int WORLD>
__global__ void __launch_bounds__(256) peer_push_kernel(PeerBufs dst, int rank, const float4* __restrict__ src,
                                                        size_t n4, size_t slot4) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  constexpr int U = 4;
  for (size_t i0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i0 < n4; i0 += U * stride) {
    float4 v[U];
#pragma unroll
    for (int u = 0; u < U; u++)
      if (i0 + u * stride < n4) v[u] = __ldcs(src + i0 + u * stride);
#pragma unroll
    for (int u = 0; u < U; u++) {
      const size_t i = i0 + u * stride;
      if (i < n4) {
#pragma unroll
        for (int p = 0; p < WORLD; p++)
          if (p != rank) dst.p[p][(size_t)rank * slot4 + i] = v[u];
      }
    }
  }
}

// dL_dsh[i] (+)= sum_v basis(normalize(mean_i - campos_v)) (x) f_v(i), f_v read from the peer that rendered view v.
// M == 16, dL_dsh 16-byte aligned (rows leave through shared memory with coalesced 128-bit stores).
template <int VB>  // views fetched per batch (independent remote loads in flight per thread)
__global__ void __launch_bounds__(PB_THREADS) sh_grad_from_peers_kernel(int P, int D, int V, PeerViews views,
                                                                        const float* __restrict__ means3D,
                                                                        float* __restrict__ dL_dsh, int accumulate) {
  extern __shared__ float4 s_tile[];  // [PB_THREADS][13]
  const int tid = threadIdx.x;
  const int base = blockIdx.x * PB_THREADS;
  const int i = base + tid;
  const int nG = min(PB_THREADS, P - base);
  const size_t I = (size_t)i;
  float v[48];
#pragma unroll
  for (int k = 0; k < 48; k++) v[k] = 0.f;
  if (i < P) {
    const float p0 = __ldg(means3D + 3 * I), p1 = __ldg(means3D + 3 * I + 1), p2 = __ldg(means3D + 3 * I + 2);
    for (int w0 = 0; w0 < V; w0 += VB) {
      float f[VB][3];
#pragma unroll
      for (int b = 0; b < VB; b++) {
        if (w0 + b < V) {
          const float* src = views.fac[w0 + b] + 3 * I;
          f[b][0] = src[0]; f[b][1] = src[1]; f[b][2] = src[2];
        } else {
          f[b][0] = f[b][1] = f[b][2] = 0.f;
        }
      }
#pragma unroll
      for (int b = 0; b < VB; b++) {
        if (w0 + b >= V || (f[b][0] == 0.f && f[b][1] == 0.f && f[b][2] == 0.f)) continue;  // invisible in this view
        const float* cam = views.fac[w0 + b] + 3 * (size_t)P;
        const float dox = p0 - cam[0], doy = p1 - cam[1], doz = p2 - cam[2];
        const float len = sqrtf(dox * dox + doy * doy + doz * doz);
        float basis[16];
#pragma unroll
        for (int k = 0; k < 16; k++) basis[k] = 0.f;
        sh_basis(D, dox / len, doy / len, doz / len, basis);
#pragma unroll
        for (int k = 0; k < 16; k++) {
          v[3 * k] = fmaf(basis[k], f[b][0], v[3 * k]);
          v[3 * k + 1] = fmaf(basis[k], f[b][1], v[3 * k + 1]);
          v[3 * k + 2] = fmaf(basis[k], f[b][2], v[3 * k + 2]);
        }
      }
    }
  }
  float4* row4 = s_tile + tid * 13;
  if (i < P) {
#pragma unroll
    for (int k = 0; k < 12; k++) row4[k] = make_float4(v[4 * k], v[4 * k + 1], v[4 * k + 2], v[4 * k + 3]);
  }
  __syncthreads();
  float4* __restrict__ dst4 = reinterpret_cast<float4*>(dL_dsh + (size_t)base * 48);
#pragma unroll 4
  for (int j = tid; j < nG * 12; j += PB_THREADS) {
    const int row = j / 12, col = j - row * 12;
    float4 q = s_tile[row * 13 + col];
    if (accumulate) {
      const float4 o = dst4[j];
      q.x += o.x; q.y += o.y; q.z += o.z; q.w += o.w;
    }
    __stcs(dst4 + j, q);
  }
}

}  // namespace gsr
