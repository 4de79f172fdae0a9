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
  const float* H;     // homographies, 9 floats each (nullptr: analytic mode below)
  long long hst[4];   // element strides of H for (n, d, y, x); 0 = broadcast
  // analytic mode: H(n, d, y, x) = AB[n][0] - AB[n][1] / (depth + 1e-9) with the plane schedule of get_homographies
  const float* AB;    // [N,2,3,3] from vis_setup_kernel
  const float* dstart;  // depth_start, element (n,y,x) at n*sst[0] + y*sst[1] + x*sst[2]
  const float* dinter;  // depth_interval, same addressing with ist
  long long sst[3], ist[3];
  int inv;
  int N, C, h, w, D, groups;
  float* out;         // fused: cost [N,G,D,h,w]; warp only: warped [N,C,D,h,w]
};

constexpr int kVisPlanes = 4;  // consecutive planes per thread: reference features and homography rows are loaded once per
                               // thread for four voxels, and four independent gathers per channel are in flight

struct VisTap {
  int o_nw, o_ne, o_sw, o_se;
  float nw, ne, sw, se;
};

template <bool kAnalytic>
__device__ __forceinline__ VisTap vis_tap(const VisArgs& A, int n, int d, int yi, int xi) {
  float Hm[9];
  if (!kAnalytic) {
    const float* Hp = A.H + n * A.hst[0] + d * A.hst[1] + yi * A.hst[2] + xi * A.hst[3];
#pragma unroll
    for (int k = 0; k < 9; ++k) Hm[k] = __ldg(Hp + k);
  } else {  // get_homographies, utils.py:113-126 (schedule) and :139-146 (H = A - B / (depth + 1e-9))
    const float ds = __ldg(A.dstart + n * A.sst[0] + yi * A.sst[1] + xi * A.sst[2]);
    const float di = __ldg(A.dinter + n * A.ist[0] + yi * A.ist[1] + xi * A.ist[2]);
    float depth;
    if (!A.inv) {
      depth = fadd(ds, fmul(di, (float)d));                                               // :114-116
    } else {
      const float dend = fadd(ds, fmul((float)(A.D - 1), di));                            // :118
      const float i0 = fdiv(1.0f, fadd(dend, 1e-9f));
      const float step = fdiv(fsub(fdiv(1.0f, fadd(ds, 1e-9f)), i0), fadd((float)(A.D - 1), 1e-9f));  // :119
      depth = fdiv(1.0f, fadd(i0, fmul(step, (float)d)));                                 // :120-123
    }
    const float r = fdiv(1.0f, fadd(depth, 1e-9f));
    const float* ab = A.AB + (size_t)n * 18;
#pragma unroll
    for (int k = 0; k < 9; ++k) Hm[k] = __fmaf_rn(-__ldg(ab + 9 + k), r, __ldg(ab + k));
  }
  const float x = (float)xi + 0.5f, y = (float)yi + 0.5f;  // get_pixel_grids, utils.py:155-160
  float q[3];
#pragma unroll
  for (int r = 0; r < 3; ++r) {  // H @ (x, y, 1), utils.py:184
    float acc = fmul(Hm[r * 3 + 0], x);
    acc = __fmaf_rn(Hm[r * 3 + 1], y, acc);
    q[r] = __fmaf_rn(Hm[r * 3 + 2], 1.0f, acc);
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

template <bool kFused, bool kAnalytic>
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
  for (int k = 0; k < kVisPlanes; ++k) tap[k] = vis_tap<kAnalytic>(A, n, min(d0 + k, A.D - 1), yi, xi);
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
  if (fused) vis_homography_kernel<true, false><<<grid, 256, 0, (cudaStream_t)stream>>>(A);
  else vis_homography_kernel<false, false><<<grid, 256, 0, (cudaStream_t)stream>>>(A);
  return (int)cudaGetLastError();
}


// Per sample: A = K_r R_r R_l^T K_l^-1 and B = K_r R_r (c_r - c_l) n^T R_l^T K_l^-1 with c = -R^T t and n^T = third row of
// R_l (get_homographies, utils.py:101-146), so that H(depth) = A - B / (depth + 1e-9).  Evaluated in double, rounded once.
// cams [N,2,4,4]: [:,0] = extrinsics [R|t], [:,1] = intrinsics.
__global__ void vis_setup_kernel(const float* __restrict__ left_cam, const float* __restrict__ right_cam, int N,
                                 float* __restrict__ AB) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  const float* L = left_cam + (size_t)n * 32;
  const float* Rc = right_cam + (size_t)n * 32;
  double Rl[9], Rr[9], tl[3], tr[3], Kl[9], Kr[9];
  for (int i = 0; i < 3; ++i) {
    for (int j = 0; j < 3; ++j) {
      Rl[i * 3 + j] = L[i * 4 + j]; Rr[i * 3 + j] = Rc[i * 4 + j];
      Kl[i * 3 + j] = L[16 + i * 4 + j]; Kr[i * 3 + j] = Rc[16 + i * 4 + j];
    }
    tl[i] = L[i * 4 + 3]; tr[i] = Rc[i * 4 + 3];
  }
  // K_l^-1 by the adjugate
  double Ki[9];
  {
    const double a = Kl[0], b = Kl[1], c = Kl[2], d = Kl[3], e = Kl[4], f = Kl[5], g = Kl[6], h = Kl[7], k = Kl[8];
    const double c00 = e * k - f * h, c01 = -(d * k - f * g), c02 = d * h - e * g;
    const double det = a * c00 + b * c01 + c * c02;
    Ki[0] = c00 / det; Ki[1] = -(b * k - c * h) / det; Ki[2] = (b * f - c * e) / det;
    Ki[3] = c01 / det; Ki[4] = (a * k - c * g) / det;  Ki[5] = -(a * f - c * d) / det;
    Ki[6] = c02 / det; Ki[7] = -(a * h - b * g) / det; Ki[8] = (a * e - b * d) / det;
  }
  auto mm = [](const double* X, const double* Y, double* Z) {
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) Z[i * 3 + j] = X[i * 3 + 0] * Y[0 * 3 + j] + X[i * 3 + 1] * Y[1 * 3 + j] + X[i * 3 + 2] * Y[2 * 3 + j];
  };
  double RlT[9], M1[9], KR[9], Am[9], outer[9], T1[9], Bm[9], crel[3];
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) RlT[i * 3 + j] = Rl[j * 3 + i];
  mm(RlT, Ki, M1);   // R_l^T K_l^-1, :132
  mm(Kr, Rr, KR);
  mm(KR, M1, Am);
  for (int i = 0; i < 3; ++i) {  // c_right - c_left, c = -R^T t (:126-128)
    const double cl = -(Rl[0 * 3 + i] * tl[0] + Rl[1 * 3 + i] * tl[1] + Rl[2 * 3 + i] * tl[2]);
    const double cr = -(Rr[0 * 3 + i] * tr[0] + Rr[1 * 3 + i] * tr[1] + Rr[2 * 3 + i] * tr[2]);
    crel[i] = cr - cl;
  }
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) outer[i * 3 + j] = crel[i] * Rl[2 * 3 + j];  // c_rel @ fronto_direction, :131
  mm(outer, M1, T1);
  mm(KR, T1, Bm);
  float* o = AB + (size_t)n * 18;
  for (int k = 0; k < 9; ++k) { o[k] = (float)Am[k]; o[9 + k] = (float)Bm[k]; }
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

extern "C" int rmvd_vis_cost_volume_fwd(const float* ref, const float* src, const float* left_cam, const float* right_cam,
                                        const float* depth_start, const int64_t* start_stride, const float* depth_interval,
                                        const int64_t* interval_stride, int N, int C, int h, int w, int D, int inv, int groups,
                                        float* cost, void* workspace, size_t workspace_bytes, void* stream) {
  using namespace rmvd;
  if (!ref || !src || !left_cam || !right_cam || !depth_start || !start_stride || !depth_interval || !interval_stride || !cost)
    return RMVD_E_NULL;
  if (N <= 0 || C <= 0 || h <= 0 || w <= 0 || D <= 0 || N > 65535 || h >= 32768 || w >= 32768) return RMVD_E_SHAPE;
  if (groups <= 0 || C % groups != 0) return RMVD_E_SHAPE;
  if (!workspace || workspace_bytes < (size_t)N * 18 * sizeof(float)) return RMVD_E_WORKSPACE;
  cudaStream_t st = (cudaStream_t)stream;
  float* AB = (float*)workspace;
  vis_setup_kernel<<<(N + 63) / 64, 64, 0, st>>>(left_cam, right_cam, N, AB);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return (int)e;
  VisArgs A{};
  A.ref = ref; A.src = src; A.H = nullptr; A.AB = AB; A.dstart = depth_start; A.dinter = depth_interval; A.inv = inv;
  for (int k = 0; k < 3; ++k) { A.sst[k] = start_stride[k]; A.ist[k] = interval_stride[k]; }
  A.N = N; A.C = C; A.h = h; A.w = w; A.D = D; A.groups = groups; A.out = cost;
  const long long tot = (long long)((D + kVisPlanes - 1) / kVisPlanes) * h * w;
  dim3 grid((unsigned)((tot + 255) / 256), N);
  vis_homography_kernel<true, true><<<grid, 256, 0, st>>>(A);
  return (int)cudaGetLastError();
}
