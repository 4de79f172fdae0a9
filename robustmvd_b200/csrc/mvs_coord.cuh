// Coordinate chains of the MVSNet-family warps (shared by mvs.cu and mvs_var.cu).
#pragma once
#include "common.cuh"

namespace rmvd {

constexpr int kMaxMvsViews = 16;  // source views per launch

// transform = src_proj @ ref_proj_inv (utils.py:234): k-ascending, every multiply and add rounded (oracle/mvsnet.py:_matmul4)
__device__ __forceinline__ float matmul4_elem(const float* a, const float* b, int i, int j) {
  float acc = fmul(a[i * 4 + 0], b[0 * 4 + j]);
#pragma unroll
  for (int k = 1; k < 4; ++k) acc = fadd(acc, fmul(a[i * 4 + k], b[k * 4 + j]));
  return acc;
}

// Source-view sampling coordinate of key pixel (xi, yi) at depth d (utils.py:240-264): returns the ATen-unnormalised
// coordinate whose floor is the corner index.  M = rows 0..2 of the 4x4 transform (R | T), row-major 3x4.
__device__ __forceinline__ void homo_coord(const float* M, int xi, int yi, float d, int H, int W, float& ix, float& iy) {
  const float gx = fmul((float)xi, d), gy = fmul((float)yi, d), gz = d;  // :248
  float p[3];
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    float acc = fmul(M[r * 4 + 0], gx);
    acc = __fmaf_rn(M[r * 4 + 1], gy, acc);
    acc = __fmaf_rn(M[r * 4 + 2], gz, acc);
    p[r] = fadd(acc, M[r * 4 + 3]);  // :251
  }
  const float sx = fdiv(p[0], p[2]), sy = fdiv(p[1], p[2]);       // :253
  const float xn = fsub(fdiv(sx, (float)(W - 1) * 0.5f), 1.0f);   // :256
  const float yn = fsub(fdiv(sy, (float)(H - 1) * 0.5f), 1.0f);   // :257
  ix = grid_unnormalize(xn, W);                                    // grid_sample, align_corners=False (:262)
  iy = grid_unnormalize(yn, H);
}

// CVP-MVSNet's chain (cvp_mvsnet_components.py:414-430, :217-233): rotate the integer pixel (x, y, 1) first, then scale by
// the (per-pixel) depth hypothesis and translate; same normalisation and grid_sample convention as above.
__device__ __forceinline__ void cvp_coord(const float* M, int xi, int yi, float d, int H, int W, float& ix, float& iy) {
  float p[3];
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    float acc = fmul(M[r * 4 + 0], (float)xi);
    acc = __fmaf_rn(M[r * 4 + 1], (float)yi, acc);
    acc = __fmaf_rn(M[r * 4 + 2], 1.0f, acc);     // rot @ xyz, :414
    p[r] = fadd(fmul(acc, d), M[r * 4 + 3]);      // * depth_hypos + trans, :416-421
  }
  const float sx = fdiv(p[0], p[2]), sy = fdiv(p[1], p[2]);       // :422
  const float xn = fsub(fdiv(sx, (float)(W - 1) * 0.5f), 1.0f);   // :423
  const float yn = fsub(fdiv(sy, (float)(H - 1) * 0.5f), 1.0f);   // :424
  ix = grid_unnormalize(xn, W);
  iy = grid_unnormalize(yn, H);
}

// Same chain with the divisions through shared reciprocals (common.cuh: bit-identical to div.rn inside the guarded range;
// operands outside it take homo_coord() itself).  y_w, y_h = rcp_refined((W - 1) / 2), rcp_refined((H - 1) / 2).
__device__ __forceinline__ void homo_coord_shared(const float* M, int xi, int yi, float d, int H, int W, float y_w, float y_h,
                                                  float& ix, float& iy) {
  const float gx = fmul((float)xi, d), gy = fmul((float)yi, d), gz = d;  // :248
  float p[3];
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    float acc = fmul(M[r * 4 + 0], gx);
    acc = __fmaf_rn(M[r * 4 + 1], gy, acc);
    acc = __fmaf_rn(M[r * 4 + 2], gz, acc);
    p[r] = fadd(acc, M[r * 4 + 3]);  // :251
  }
  // numerators and quotients: zero or inside [2^-60, 2^60], so that no quotient or residual is subnormal
  auto in_range = [](float v) { const float a = fabsf(v); return (a == 0.0f) || ((a >= kDivLo) && (a <= kDivHi)); };
  const float az = fabsf(p[2]);
  bool fast = (az >= kDivLo) && (az <= kDivHi) && in_range(p[0]) && in_range(p[1]) && (W > 1) && (H > 1);
  const float y = rcp_refined(p[2]);
  const float sx = div_shared(p[0], p[2], y), sy = div_shared(p[1], p[2], y);
  fast = fast && in_range(sx) && in_range(sy);
  if (fast) {
    const float xn = fsub(div_shared(sx, (float)(W - 1) * 0.5f, y_w), 1.0f);
    const float yn = fsub(div_shared(sy, (float)(H - 1) * 0.5f, y_h), 1.0f);
    ix = grid_unnormalize(xn, W);
    iy = grid_unnormalize(yn, H);
  } else {
    homo_coord(M, xi, yi, d, H, W, ix, iy);
  }
}

}  // namespace rmvd
