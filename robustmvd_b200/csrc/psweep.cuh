// Argument block shared by the plane-sweep correlation kernels (psweep.cu: general direct kernel,
// psweep_band.cu: banded-GEMM kernel).
#pragma once
#include "common.cuh"

namespace rmvd {

constexpr int kMaxViews = 16;   // source views per launch (more views: the host loops)
constexpr int kTilePx = 32;     // key pixels per CTA
constexpr int kWarps = 8;       // warps per CTA; each owns kTilePx / kWarps pixels
constexpr int kPxPerWarp = kTilePx / kWarps;

struct PsweepArgs {
  const float* key;                  // [N,C,h,w]
  const float* src[kMaxViews];       // channels-last copies [N,hs*ws,C]
  const float* Kk;                   // [N,3,3]
  const float* Ks[kMaxViews];        // [N,3,3]
  const float* T[kMaxViews];         // [N,4,4]
  const float* invd;
  long long ist[4];                  // element strides of invd for (n,s,y,x)
  const float* grids;                // kFromGrid: [N,S,2,h,w] sampling positions in source pixels
  const uint8_t* grid_mask;          // kFromGrid: optional [N,S,h,w]
  int hs[kMaxViews], ws[kMaxViews];
  int N, C, h, w, S, s_begin, s_end, norm, out_S, out_s0, v0;
  float* corr;                       // [V,N,out_S,h,w]
  float* mask;
  int32_t* dbg_ix0;
  int32_t* dbg_iy0;
  uint8_t* dbg_geom;
};

// psweep_band.cu: launches the banded kernel when the shape qualifies; returns -1000 when it does not.
int launch_psweep_band(const PsweepArgs& A, const float* key_nhwc, int nviews, cudaStream_t st);

}  // namespace rmvd
