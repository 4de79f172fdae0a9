// Backward of the MVSNet cost-volume path (training direction of mvsnet.py:123-141; the reference gets it from autograd through
// homo_warp's grid_sample, the variance expression and softmax + depth_regression).  Two kernels:
//
//   mvs_variance_bwd_kernel   d variance / d feats.  With x_0 = ref, x_v = warp_v(src_v) and mean = (sum_v x_v) / V the variance is
//                             (sum x_v^2) / V - mean^2 (:135), so d var / d x_v = (2 / V) (x_v - mean): the key view receives it summed
//                             over the planes, a source view through the transpose of its bilinear gather (grid_sample's zeros
//                             padding: corners outside the map carry no gradient).  The warped volumes are recomputed (two passes over
//                             the views per voxel: mean first, then the scatter) -- nothing of the forward is saved.  Sampling
//                             positions carry no gradient (depth hypotheses and projections are data).
//   mvs_regress_bwd_kernel    d depth / d cost_reg for depth = sum_d softmax(cost)_d * depth_d (:140-141):
//                             grad_cost_d = g * p_d * (depth_d - depth).  The confidence is computed under no_grad (:143-158).
//
// Direct forms (thread = voxel x 8 channels, float atomics into the feature gradients): the training direction is a "next" row of
// SURVEY.md 8(f) and is built for correctness first; see DESIGN.md section 4.2 for its measured cost.
#include "mvs_coord.cuh"

#include "../../include/rmvd_b200.h"

namespace rmvd {
namespace {

struct Corner4 {
  int o[4];
  float wgt[4];
};

__device__ __forceinline__ Corner4 corners_of(float ix, float iy, int H, int W) {
  const Bilinear b = bilinear_setup(ix, iy, H, W);
  const int x0 = isnan(b.x0f) ? -2 : sat_index(b.x0f, W), y0 = isnan(b.y0f) ? -2 : sat_index(b.y0f, H);
  const int x0c = clampi(x0, 0, W - 1), x1c = clampi(x0 + 1, 0, W - 1);
  const int y0c = clampi(y0, 0, H - 1), y1c = clampi(y0 + 1, 0, H - 1);
  Corner4 c;
  c.o[0] = y0c * W + x0c; c.o[1] = y0c * W + x1c; c.o[2] = y1c * W + x0c; c.o[3] = y1c * W + x1c;
  c.wgt[0] = b.nw; c.wgt[1] = b.ne; c.wgt[2] = b.sw; c.wgt[3] = b.se;      // 0 for corners outside the map
  return c;
}

__device__ __forceinline__ float warp_val(const float* __restrict__ img, const Corner4& c) {
  float acc = fmul(__ldg(img + c.o[0]), c.wgt[0]);
#pragma unroll
  for (int k = 1; k < 4; ++k) acc = fadd(acc, fmul(__ldg(img + c.o[k]), c.wgt[k]));
  return acc;
}

constexpr int kBwdCh = 8;

__global__ void __launch_bounds__(256) mvs_variance_bwd_kernel(const float* __restrict__ feats, const float* __restrict__ proj,
                                                               const float* __restrict__ depths, const float* __restrict__ grad_var,
                                                               int B, int Vt, int C, int h, int w, int D, float* __restrict__ grad_feats) {
  __shared__ float s_M[kMaxMvsViews][12];
  const int b = blockIdx.z;
  const int P = h * w;
  const int c0 = blockIdx.y * kBwdCh;
  for (int e = threadIdx.x; e < (Vt - 1) * 12; e += blockDim.x) {
    const int v = 1 + e / 12, i = (e % 12) / 4, j = e % 4;
    s_M[v - 1][i * 4 + j] = matmul4_elem(proj + ((size_t)b * Vt + v) * 16, proj + ((size_t)b * Vt) * 16, i, j);
  }
  __syncthreads();
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (long long)D * P) return;
  const int d = (int)(idx / P), p = (int)(idx - (long long)d * P);
  const int yi = p / w, xi = p - yi * w;
  const float depth = depths[(size_t)b * D + d];
  const float inv_v = 1.0f / (float)Vt;

  float g[kBwdCh], mean[kBwdCh];
  bool any = false;
  const float* ref = feats + ((size_t)b * Vt) * C * P + p;
#pragma unroll
  for (int k = 0; k < kBwdCh; ++k) {
    const bool in = c0 + k < C;
    g[k] = in ? __ldg(grad_var + (((size_t)b * C + c0 + k) * D + d) * P + p) : 0.f;
    mean[k] = in ? __ldg(ref + (size_t)(c0 + k) * P) : 0.f;
    any = any || (g[k] != 0.f);
  }
  if (!any) return;
  // pass 1: mean over the views
  for (int v = 1; v < Vt; ++v) {
    float ix, iy;
    homo_coord(s_M[v - 1], xi, yi, depth, h, w, ix, iy);
    const Corner4 c = corners_of(ix, iy, h, w);
    const float* src = feats + ((size_t)b * Vt + v) * C * P;
#pragma unroll
    for (int k = 0; k < kBwdCh; ++k)
      if (c0 + k < C) mean[k] += warp_val(src + (size_t)(c0 + k) * P, c);
  }
#pragma unroll
  for (int k = 0; k < kBwdCh; ++k) mean[k] *= inv_v;
  // key view: (2 / V) g (ref - mean), summed over the planes by atomics
  float* gref = grad_feats + ((size_t)b * Vt) * C * P + p;
#pragma unroll
  for (int k = 0; k < kBwdCh; ++k) {
    if (c0 + k < C && g[k] != 0.f) {
      const float r = __ldg(ref + (size_t)(c0 + k) * P);
      atomicAdd(gref + (size_t)(c0 + k) * P, 2.0f * inv_v * g[k] * (r - mean[k]));
    }
  }
  // pass 2: source views, transpose of the gather
  for (int v = 1; v < Vt; ++v) {
    float ix, iy;
    homo_coord(s_M[v - 1], xi, yi, depth, h, w, ix, iy);
    const Corner4 c = corners_of(ix, iy, h, w);
    if (c.wgt[0] == 0.f && c.wgt[1] == 0.f && c.wgt[2] == 0.f && c.wgt[3] == 0.f) continue;   // sample entirely outside (x_v = 0: only the mean terms above see it)
    const float* src = feats + ((size_t)b * Vt + v) * C * P;
    float* gsrc = grad_feats + ((size_t)b * Vt + v) * C * P;
#pragma unroll
    for (int k = 0; k < kBwdCh; ++k) {
      if (c0 + k < C && g[k] != 0.f) {
        const float x = warp_val(src + (size_t)(c0 + k) * P, c);
        const float gx = 2.0f * inv_v * g[k] * (x - mean[k]);
#pragma unroll
        for (int q = 0; q < 4; ++q)
          if (c.wgt[q] != 0.f) atomicAdd(gsrc + (size_t)(c0 + k) * P + c.o[q], gx * c.wgt[q]);
      }
    }
  }
}

// thread = pixel; sweep 1: running maximum and normaliser + expected depth, sweep 2: the gradient row by row (128-byte rows)
__global__ void __launch_bounds__(128) mvs_regress_bwd_kernel(const float* __restrict__ cost, const float* __restrict__ depths,
                                                              const float* __restrict__ grad_depth, int D, int P, float* __restrict__ grad_cost) {
  const int b = blockIdx.y;
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= P) return;
  const float* x = cost + (size_t)b * D * P + p;
  const float* dv = depths + (size_t)b * D;
  float m = -INFINITY, Z = 0.f, A = 0.f;
  for (int d = 0; d < D; ++d) {
    const float v = __ldg(x + (size_t)d * P);
    if (v > m) { const float r = expf(m - v); Z *= r; A *= r; m = v; }
    const float e = expf(v - m);
    Z += e;
    A = fmaf(e, __ldg(dv + d), A);
  }
  const float depth = A / Z, gz = __ldg(grad_depth + (size_t)b * P + p) / Z;
  float* go = grad_cost + (size_t)b * D * P + p;
  for (int d = 0; d < D; ++d) {
    const float e = expf(__ldg(x + (size_t)d * P) - m);
    go[(size_t)d * P] = gz * e * (__ldg(dv + d) - depth);
  }
}

}  // namespace
}  // namespace rmvd

extern "C" int rmvd_mvs_variance_bwd(const float* feats, const float* proj_mats, const float* depth_values, const float* grad_variance,
                                     int B, int Vt, int C, int h, int w, int D, float* grad_feats, void* stream) {
  if (!feats || !proj_mats || !depth_values || !grad_variance || !grad_feats) return RMVD_E_NULL;
  if (B <= 0 || Vt < 1 || Vt - 1 > rmvd::kMaxMvsViews || C <= 0 || h <= 0 || w <= 0 || D <= 0 || B > 65535) return RMVD_E_SHAPE;
  cudaStream_t st = (cudaStream_t)stream;
  const size_t n = (size_t)B * Vt * C * h * w;
  cudaError_t e = cudaMemsetAsync(grad_feats, 0, n * sizeof(float), st);
  if (e != cudaSuccess) return (int)e;
  const long long vox = (long long)D * h * w;
  const long long blocks = (vox + 255) / 256;
  if (blocks > 0x7fffffffLL || (C + rmvd::kBwdCh - 1) / rmvd::kBwdCh > 65535) return RMVD_E_SHAPE;
  dim3 grid((unsigned)blocks, (C + rmvd::kBwdCh - 1) / rmvd::kBwdCh, B);
  rmvd::mvs_variance_bwd_kernel<<<grid, 256, 0, st>>>(feats, proj_mats, depth_values, grad_variance, B, Vt, C, h, w, D, grad_feats);
  return (int)cudaGetLastError();
}

extern "C" int rmvd_mvs_regress_bwd(const float* cost, const float* depth_values, const float* grad_depth, int B, int D, int h, int w,
                                    float* grad_cost, void* stream) {
  if (!cost || !depth_values || !grad_depth || !grad_cost) return RMVD_E_NULL;
  if (B <= 0 || D <= 0 || h <= 0 || w <= 0 || B > 65535) return RMVD_E_SHAPE;
  const int P = h * w;
  dim3 grid((P + 127) / 128, B);
  rmvd::mvs_regress_bwd_kernel<<<grid, 128, 0, (cudaStream_t)stream>>>(cost, depth_values, grad_depth, D, P, grad_cost);
  return (int)cudaGetLastError();
}
