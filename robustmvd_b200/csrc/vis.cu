// Vis-MVSNet cost volume for sm_100a (first part): homography warping of the source features to every
// depth hypothesis (blocks/utils.py:179-189 `homography_warping`, :164-176 `interpolate`, :155-160 `get_pixel_grids`) fused
// with the group-wise correlation against the reference features (blocks/utils.py:71-89 `groupwise_correlation`, call site
// vis_mvsnet_singlestage.py:230-242), so that the warped volume [n,c,d,h,w] of `build_cost_volume`
// (vis_mvsnet_singlestage.py:86-122) is never written:
//   cost[n,g,d,y,x] = sum_{c in group g} ref[n,c,y,x] * bilinear(src[n,c], u, v),
//   (hx,hy,hz) = H[n,d,(y,x)] . (x + 0.5, y + 0.5, 1),  (u,v) = (hx,hy) / (hz + 1e-9),
//   grid = clamp(2 * (u / W, v / H) - 1, -1.1, 1.1), grid_sample(bilinear, zeros, align_corners=False).
// Thread = four consecutive planes of one pixel, x fastest across lanes: reference features and outputs are 128-byte rows per warp, source texels of adjacent
// lanes are adjacent.  The homographies come from the host-side mirror of `get_homographies` (per plane, or per pixel for
// the refinement stages); element strides select the broadcast form.
#include "common.cuh"

namespace rmvd {

struct VisArgs {
  const float* ref;   // [N,C,h,w] (fused mode)
  const float* src;   // [N,C,h,w]
  const float* H;     // homographies, 9 floats each
  long long hst[4];   // element strides of H for (n, d, y, x); 0 = broadcast
  int N, C, h, w, D, groups;
  float* out;         // fused: cost [N,G,D,h,w]; warp only: warped [N,C,D,h,w]
};

constexpr int kVisPlanes = 4;  // consecutive planes per thread: reference features and homography rows are loaded once per
                               // thread for four voxels, and four independent gathers per channel are in flight

struct VisTap {
  int o_nw, o_ne, o_sw, o_se;
  float nw, ne, sw, se;
};

__device__ __forceinline__ VisTap vis_tap(const VisArgs& A, int n, int d, int yi, int xi) {
  const float* Hm = A.H + n * A.hst[0] + d * A.hst[1] + yi * A.hst[2] + xi * A.hst[3];
  const float x = (float)xi + 0.5f, y = (float)yi + 0.5f;  // get_pixel_grids, utils.py:155-160
  float q[3];
#pragma unroll
  for (int r = 0; r < 3; ++r) {  // H @ (x, y, 1), utils.py:184
    float acc = fmul(__ldg(Hm + r * 3 + 0), x);
    acc = __fmaf_rn(__ldg(Hm + r * 3 + 1), y, acc);
    q[r] = __fmaf_rn(__ldg(Hm + r * 3 + 2), 1.0f, acc);
  }
  const float den = fadd(q[2], 1e-9f);                       // :185
  float gx = fdiv(fdiv(q[0], den), (float)A.w);              // :185, :168
  float gy = fdiv(fdiv(q[1], den), (float)A.h);              // :169
  gx = fsub(fmul(gx, 2.0f), 1.0f);                           // :170
  gy = fsub(fmul(gy, 2.0f), 1.0f);
  gx = fminf(fmaxf(gx, -1.1f), 1.1f);                        // clamp(-1.1, 1.1); a NaN coordinate samples nothing here
  gy = fminf(fmaxf(gy, -1.1f), 1.1f);                        // (the reference raises NanError, :173)
  const Bilinear b = bilinear_setup(grid_unnormalize(gx, A.w), grid_unnormalize(gy, A.h), A.h, A.w);
  const int x0 = isnan(b.x0f) ? -2 : sat_index(b.x0f, A.w), y0 = isnan(b.y0f) ? -2 : sat_index(b.y0f, A.h);
  const int x0c = clampi(x0, 0, A.w - 1), x1c = clampi(x0 + 1, 0, A.w - 1);
  const int y0c = clampi(y0, 0, A.h - 1), y1c = clampi(y0 + 1, 0, A.h - 1);
  VisTap t;
  t.o_nw = y0c * A.w + x0c; t.o_ne = y0c * A.w + x1c; t.o_sw = y1c * A.w + x0c; t.o_se = y1c * A.w + x1c;
  t.nw = b.nw; t.ne = b.ne; t.sw = b.sw; t.se = b.se;
  return t;
}
__device__ __forceinline__ float vis_gather(const float* __restrict__ sc, const VisTap& t) {
  return __ldg(sc + t.o_nw) * t.nw + __ldg(sc + t.o_ne) * t.ne + __ldg(sc + t.o_sw) * t.sw + __ldg(sc + t.o_se) * t.se;
}

template <bool kFused>
__global__ void __launch_bounds__(256) vis_homography_kernel(const VisArgs A) {
  const int n = blockIdx.y;
  const int P = A.h * A.w;
  const int Dq = (A.D + kVisPlanes - 1) / kVisPlanes;
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (long long)Dq * P) return;
  const int d0 = (int)(idx / P) * kVisPlanes, p = (int)(idx % P);
  const int yi = p / A.w, xi = p - yi * A.w;
  VisTap tap[kVisPlanes];
#pragma unroll
  for (int k = 0; k < kVisPlanes; ++k) tap[k] = vis_tap(A, n, min(d0 + k, A.D - 1), yi, xi);
  const float* src = A.src + (size_t)n * A.C * P;
  const size_t DP = (size_t)A.D * P;
  if (kFused) {
    const float* ref = A.ref + (size_t)n * A.C * P + p;
    const int cpg = A.C / A.groups;
    float* out = A.out + (((size_t)n * A.groups) * A.D + d0) * P + p;
    for (int g = 0; g < A.groups; ++g) {
      float acc[kVisPlanes];
#pragma unroll
      for (int k = 0; k < kVisPlanes; ++k) acc[k] = 0.f;
      for (int cc = 0; cc < cpg; ++cc) {
        const int c = g * cpg + cc;
        const float* sc = src + (size_t)c * P;
        const float rv = __ldg(ref + (size_t)c * P);
#pragma unroll
        for (int k = 0; k < kVisPlanes; ++k) acc[k] = fmaf(rv, vis_gather(sc, tap[k]), acc[k]);
      }
#pragma unroll
      for (int k = 0; k < kVisPlanes; ++k)
        if (d0 + k < A.D) __stcs(out + (size_t)g * DP + (size_t)k * P, acc[k]);
    }
  } else {
    float* out = A.out + (((size_t)n * A.C) * A.D + d0) * P + p;
    for (int c = 0; c < A.C; ++c) {
      const float* sc = src + (size_t)c * P;
#pragma unroll
      for (int k = 0; k < kVisPlanes; ++k)
        if (d0 + k < A.D) __stcs(out + (size_t)c * DP + (size_t)k * P, vis_gather(sc, tap[k]));
    }
  }
}

static int vis_launch(bool fused, const float* ref, const float* src, const float* H, const int64_t* hst, int N, int C, int h,
                      int w, int D, int groups, float* out, void* stream) {
  if (!src || !H || !hst || !out || (fused && !ref)) return RMVD_E_NULL;
  if (N <= 0 || C <= 0 || h <= 0 || w <= 0 || D <= 0 || N > 65535 || h >= 32768 || w >= 32768) return RMVD_E_SHAPE;
  if (fused && (groups <= 0 || C % groups != 0)) return RMVD_E_SHAPE;
  VisArgs A{};
  A.ref = ref; A.src = src; A.H = H;
  for (int k = 0; k < 4; ++k) A.hst[k] = hst[k];
  A.N = N; A.C = C; A.h = h; A.w = w; A.D = D; A.groups = groups; A.out = out;
  const long long tot = (long long)((D + kVisPlanes - 1) / kVisPlanes) * h * w;
  dim3 grid((unsigned)((tot + 255) / 256), N);
  if (fused) vis_homography_kernel<true><<<grid, 256, 0, (cudaStream_t)stream>>>(A);
  else vis_homography_kernel<false><<<grid, 256, 0, (cudaStream_t)stream>>>(A);
  return (int)cudaGetLastError();
}

}  // namespace rmvd

extern "C" int rmvd_homography_groupcorr_fwd(const float* ref, const float* src, const float* H, const int64_t* h_stride,
                                             int N, int C, int h, int w, int D, int groups, float* cost, void* stream) {
  return rmvd::vis_launch(true, ref, src, H, h_stride, N, C, h, w, D, groups, cost, stream);
}

extern "C" int rmvd_homography_warp_fwd(const float* src, const float* H, const int64_t* h_stride, int N, int C, int h, int w,
                                        int D, float* warped, void* stream) {
  return rmvd::vis_launch(false, nullptr, src, H, h_stride, N, C, h, w, D, 1, warped, stream);
}
