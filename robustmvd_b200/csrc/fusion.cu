// Learned view fusion, the elementwise part (rmvd/models/blocks/learned_fusion.py:33-47): softmax of the per-view logits over
// the views (+ 1e-9), mask weighting, weighted average of the per-view correlation volumes and the fused 0/1 mask -- one pass
// over corrs / masks instead of the reference's ~10 full-volume passes (two stacks, products, sums, the division, the mask).
// the consumer immediately behind the plane-sweep correlation.  The 3x3 / 1x1 convolutions
// that produce the logits stay in cuDNN (models/nets.py).
//
// Thread = one key pixel x kFusePl consecutive planes: adjacent lanes = adjacent pixels, every load and store of a warp is
// one 128-byte row.  HBM-bound: (2 V + 2) * 4 bytes per fused voxel.
#include <cuda_runtime.h>
#include <cstdint>

#include "../../include/rmvd_b200.h"
#include "common.cuh"

namespace rmvd {

constexpr int kFuseMaxViews = 64;
constexpr int kFusePl = 8;        // planes per thread (2 * V * kFusePl independent loads in flight)
constexpr int kFuseThreads = 128;

struct FusionArgs {
  const float* logit[kFuseMaxViews];   // [N,1,h,w] each
  const float* corr[kFuseMaxViews];    // [N,S,h,w] each
  const float* mask[kFuseMaxViews];
};

__global__ void __launch_bounds__(kFuseThreads) view_fusion_kernel(const FusionArgs A, int V, int N, int S, int P,
                                                                   float* __restrict__ weights, float* __restrict__ fused_corr,
                                                                   float* __restrict__ fused_mask) {
  const int p = blockIdx.x * kFuseThreads + threadIdx.x;
  const int n = blockIdx.z;
  const int s0 = blockIdx.y * kFusePl;
  if (p >= P) return;
  // softmax over the views of this pixel's logits (:33): exp(x - max) / sum, then + 1e-9 (:33)
  float mx = -INFINITY;
  for (int v = 0; v < V; ++v) mx = fmaxf(mx, __ldg(A.logit[v] + (size_t)n * P + p));
  float den = 0.f;
  for (int v = 0; v < V; ++v) den = fadd(den, expf(fsub(__ldg(A.logit[v] + (size_t)n * P + p), mx)));
  float wsum[kFusePl], csum[kFusePl];
#pragma unroll
  for (int k = 0; k < kFusePl; ++k) { wsum[k] = 0.f; csum[k] = 0.f; }
  const size_t row = ((size_t)n * S + s0) * P + p;
  for (int v = 0; v < V; ++v) {
    const float w = fadd(fdiv(expf(fsub(__ldg(A.logit[v] + (size_t)n * P + p), mx)), den), 1e-9f);
    if (weights && blockIdx.y == 0) weights[((size_t)v * N + n) * P + p] = w;
    const float* c = A.corr[v] + row;
    const float* m = A.mask[v] + row;
    float cv[kFusePl], mv[kFusePl];
#pragma unroll
    for (int k = 0; k < kFusePl; ++k) {
      const bool in = s0 + k < S;
      cv[k] = in ? __ldcs(c + (size_t)k * P) : 0.f;     // streamed: each element is read exactly once
      mv[k] = in ? __ldcs(m + (size_t)k * P) : 0.f;
    }
#pragma unroll
    for (int k = 0; k < kFusePl; ++k) {
      const float wm = fmul(w, mv[k]);                    // view_weight * mask (:38-41)
      wsum[k] = fadd(wsum[k], wm);                        // .sum(0) over the views, in view order (:42)
      csum[k] = fadd(csum[k], fmul(cv[k], wm));           // (corrs * view_weights).sum(0) (:45-47)
    }
  }
#pragma unroll
  for (int k = 0; k < kFusePl; ++k) {
    if (s0 + k < S) {
      const float fm = (wsum[k] != 0.f) ? 1.f : 0.f;      // :43
      __stcs(fused_mask + row + (size_t)k * P, fm);
      __stcs(fused_corr + row + (size_t)k * P, fmul(fdiv(csum[k], fadd(wsum[k], 1e-9f)), fm));   // :48
    }
  }
}

}  // namespace rmvd

extern "C" int rmvd_view_fusion_fwd(const float* const* logits, const float* const* corrs, const float* const* masks, int V,
                                    int N, int S, int h, int w, float* weights, float* fused_corr, float* fused_mask,
                                    void* stream) {
  if (!logits || !corrs || !masks || !fused_corr || !fused_mask) return RMVD_E_NULL;
  if (V < 1 || V > rmvd::kFuseMaxViews || N < 1 || S < 1 || h < 1 || w < 1 || N > 65535) return RMVD_E_SHAPE;
  rmvd::FusionArgs A;
  for (int v = 0; v < rmvd::kFuseMaxViews; ++v) {
    const int u = v < V ? v : 0;
    if (!logits[u] || !corrs[u] || !masks[u]) return RMVD_E_NULL;
    A.logit[v] = logits[u]; A.corr[v] = corrs[u]; A.mask[v] = masks[u];
  }
  const int P = h * w;
  const long long ny = (S + rmvd::kFusePl - 1) / rmvd::kFusePl;
  if (ny > 65535) return RMVD_E_SHAPE;
  dim3 grid((P + rmvd::kFuseThreads - 1) / rmvd::kFuseThreads, (unsigned)ny, N);
  rmvd::view_fusion_kernel<<<grid, rmvd::kFuseThreads, 0, (cudaStream_t)stream>>>(A, V, N, S, P, weights, fused_corr, fused_mask);
  return (int)cudaGetLastError();
}
