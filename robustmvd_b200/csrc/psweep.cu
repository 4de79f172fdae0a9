// robust_mvd plane-sweep correlation for sm_100a: fused epipolar geometry + bilinear gather + channel dot.
//
// Replaces PlanesweepCorrelation.forward (planesweep_corr.py:396-521) and its helpers.  The reference's
// all-pairs (h*w) x (hs*ws) correlation matrix (:174) is never formed: bilinear interpolation is linear, so
//   corr[n,s,y,x] = mask * norm * < key[n,:,y,x], bilinear(src[n], u_s(y,x), v_s(y,x)) >.
//
// Data layout in HBM
//   key features   NCHW as given; a CTA transposes its 32-pixel tile into shared memory once.
//   source features repacked once per call to channels-last [N, hs*ws, C] in the caller's workspace, so one
//                  source texel is one contiguous C-vector (1 KB at C = 256) fetched with 128-bit loads.
//   corr, mask     [V, N, S, h, w] written once, 128-byte coalesced rows staged through shared memory.
#include <cstdlib>
#include <cstring>
#include <mutex>

#include "psweep.cuh"

namespace rmvd {

// ---------------------------------------------------------------------------------------------------------
// NCHW -> channels-last repack of one source feature map, optionally L2-normalised along C
// (normalize(x, dim=1), planesweep_corr.py:8-10, used by TorchCorr :170-172 and WarpOnlyCorr :128-129).
// One CTA: 32 texels x all channels.
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) nchw_to_nhwc_kernel(const float* __restrict__ src, float* __restrict__ dst,
                                                           int C, int P, int l2norm) {
  __shared__ float tile[32][33];
  __shared__ float s_norm[32];
  const int n = blockIdx.y;
  const int p0 = blockIdx.x * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // ty in [0,8)
  const float* s = src + (size_t)n * C * P;
  float* d = dst + (size_t)n * P * C;
  if (l2norm) {
    float acc = 0.f;
    for (int c = ty; c < C; c += 8) {
      const float v = (p0 + tx < P) ? s[(size_t)c * P + p0 + tx] : 0.f;
      acc = fmaf(v, v, acc);
    }
    tile[ty][tx] = acc;
    __syncthreads();
    if (ty == 0) {
      float t = 0.f;
#pragma unroll
      for (int k = 0; k < 8; ++k) t += tile[k][tx];
      s_norm[tx] = sqrtf(t) + 1e-9f;
    }
    __syncthreads();
  }
  for (int c0 = 0; c0 < C; c0 += 32) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int c = c0 + ty + 8 * k;
      tile[ty + 8 * k][tx] = (c < C && p0 + tx < P) ? s[(size_t)c * P + p0 + tx] : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int p = p0 + ty + 8 * k, c = c0 + tx;
      if (p < P && c < C) {
        float v = tile[tx][ty + 8 * k];
        if (l2norm) v = fdiv(v, s_norm[ty + 8 * k]);
        d[(size_t)p * C + c] = v;
      }
    }
    __syncthreads();
  }
}

// ---------------------------------------------------------------------------------------------------------
// General direct kernel.  CTA = 32 consecutive key pixels (flattened y*w+x) of one (view, sample); warp = 4
// of those pixels.  For a chunk of 32 planes, lane j derives plane j's sampling point, index and mask for the
// warp's current pixel (bit-exact chain), then the warp walks the VALID planes only: lanes split the channel
// axis, fetch the four corner texels with 128-bit loads and reduce the weighted dot product.
//   kC4 > 0 : C == 128 * kC4, key vector kept in registers; kC4 == 0 : any C, scalar channel loop.
//   kFromGrid: sampling positions come from a grids tensor (TorchCorr.forward on caller grids).
// ---------------------------------------------------------------------------------------------------------
template <int kC4, bool kFromGrid>
__global__ void __launch_bounds__(kWarps * 32) psweep_corr_kernel(const PsweepArgs A) {
  extern __shared__ __align__(16) float smem[];
  const int C = A.C, Cp = C + 4;
  float* s_ref = smem;                      // [32][Cp]
  float* s_corr = s_ref + kTilePx * Cp;     // [32 planes][33]
  float* s_mask = s_corr + 32 * 33;         // [32 planes][33]

  const int v = blockIdx.z, n = blockIdx.y;
  const int P = A.h * A.w;
  const int p0 = blockIdx.x * kTilePx;
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int hs = A.hs[v], ws = A.ws[v];
  const float* __restrict__ src = A.src[v] + (size_t)n * hs * ws * C;

  // key tile -> shared memory, transposed to [pixel][channel]
  {
    const float* key = A.key + (size_t)n * C * P;
    const int px = t & 31;
    for (int c = t >> 5; c < C; c += kWarps) s_ref[px * Cp + c] = (p0 + px < P) ? key[(size_t)c * P + p0 + px] : 0.f;
  }
  __syncthreads();
  if (A.norm == RMVD_NORM_BEFORE) {  // normalize(feat_ref, dim=1), :170-171
    for (int q = 0; q < kPxPerWarp; ++q) {
      float* r = s_ref + (warp * kPxPerWarp + q) * Cp;
      float acc = 0.f;
      for (int c = lane; c < C; c += 32) acc = fmaf(r[c], r[c], acc);
      const float nrm = sqrtf(warp_sum(acc)) + 1e-9f;
      for (int c = lane; c < C; c += 32) r[c] = fdiv(r[c], nrm);
    }
    __syncwarp();
  }

  Epi E;
  if (!kFromGrid) E = epi_from_calib(A.Kk + n * 9, A.Ks[v] + n * 9, A.T[v] + n * 16, A.h, A.w, hs, ws);
  const float scale = (A.norm == RMVD_NORM_DIM) ? sqrtf((float)C) : 1.0f;  // corr / sqrt(C), :178-179

  const size_t out_base = ((size_t)(A.v0 + v) * A.N + n) * A.out_S;

  for (int sc = A.s_begin; sc < A.s_end; sc += 32) {
    for (int q = 0; q < kPxPerWarp; ++q) {
      const int pl = warp * kPxPerWarp + q;
      const int p = p0 + pl;
      const int s = sc + lane;
      bool valid = false;
      int pack = 0;
      float wnw = 0.f, wne = 0.f, wsw = 0.f, wse = 0.f;
      if (p < P && s < A.s_end) {
        const int yi = p / A.w, xi = p - yi * A.w;
        float us, vs;
        bool geom;
        if (kFromGrid) {
          const size_t gi = (((size_t)n * A.S + s) * 2) * P + p;
          us = A.grids[gi];
          vs = A.grids[gi + P];
          geom = A.grid_mask ? (A.grid_mask[((size_t)n * A.S + s) * P + p] != 0) : true;
        } else {
          const EpiPix ep = epi_pixel(E, xi, yi);
          const float d = A.invd[n * A.ist[0] + s * A.ist[1] + yi * A.ist[2] + xi * A.ist[3]];
          epi_sample(E, ep, d, us, vs, geom);
        }
        float ix = pix_to_unnorm(us, ws), iy = pix_to_unnorm(vs, hs);
        if (kFromGrid && A.border) {   // ATen clip_coordinates with the CPU's std::min / std::max: nan -> size - 1
          ix = isnan(ix) ? (float)(ws - 1) : fminf((float)(ws - 1), fmaxf(ix, 0.f));
          iy = isnan(iy) ? (float)(hs - 1) : fminf((float)(hs - 1), fmaxf(iy, 0.f));
        }
        const Bilinear b = bilinear_setup(ix, iy, hs, ws);
        valid = ((kFromGrid && A.border) || sampling_mask(b.msum)) && geom;
        wnw = b.nw; wne = b.ne; wsw = b.sw; wse = b.se;
        const int x0 = sat_index(b.x0f, ws), y0 = sat_index(b.y0f, hs);
        const int x0c = clampi(x0, 0, ws - 1), x1c = clampi(x0 + 1, 0, ws - 1);
        const int y0c = clampi(y0, 0, hs - 1), y1c = clampi(y0 + 1, 0, hs - 1);
        pack = x0c | (y0c << 15) | ((x1c - x0c) << 30) | ((y1c - y0c) << 31);
        if (A.dbg_ix0) {
          const size_t o = (out_base + A.out_s0 + (s - A.s_begin)) * P + p;
          A.dbg_ix0[o] = x0;
          A.dbg_iy0[o] = y0;
          if (A.dbg_geom) A.dbg_geom[o] = geom ? 1 : 0;
        }
      }
      unsigned vm = __ballot_sync(kFullMask, valid);
      float res = 0.f;

      float4 r[kC4 > 0 ? kC4 : 1];
      if (kC4 > 0) {
#pragma unroll
        for (int k = 0; k < kC4; ++k) r[k] = *reinterpret_cast<const float4*>(s_ref + pl * Cp + k * 128 + lane * 4);
      }
      while (vm) {
        const int j = __ffs(vm) - 1;
        vm &= vm - 1;
        const int pk = __shfl_sync(kFullMask, pack, j);
        const float nw = __shfl_sync(kFullMask, wnw, j), ne = __shfl_sync(kFullMask, wne, j);
        const float sw = __shfl_sync(kFullMask, wsw, j), se = __shfl_sync(kFullMask, wse, j);
        const int x0c = pk & 0x7fff, y0c = (pk >> 15) & 0x7fff;
        const int dx = (pk >> 30) & 1, dy = (pk >> 31) & 1;
        const float* t_nw = src + ((size_t)y0c * ws + x0c) * C;
        const float* t_ne = t_nw + dx * C;
        const float* t_sw = t_nw + (size_t)dy * ws * C;
        const float* t_se = t_sw + dx * C;
        float a_nw = 0.f, a_ne = 0.f, a_sw = 0.f, a_se = 0.f;
        if (kC4 > 0) {
#pragma unroll
          for (int k = 0; k < kC4; ++k) {
            const int o = k * 128 + lane * 4;
            const float4 q0 = ldg4(t_nw + o), q1 = ldg4(t_ne + o), q2 = ldg4(t_sw + o), q3 = ldg4(t_se + o);
            a_nw = dot4(r[k], q0, a_nw);
            a_ne = dot4(r[k], q1, a_ne);
            a_sw = dot4(r[k], q2, a_sw);
            a_se = dot4(r[k], q3, a_se);
          }
        } else {
          const float* rr = s_ref + pl * Cp;
          for (int c = lane; c < C; c += 32) {
            const float rv = rr[c];
            a_nw = fmaf(rv, __ldg(t_nw + c), a_nw);
            a_ne = fmaf(rv, __ldg(t_ne + c), a_ne);
            a_sw = fmaf(rv, __ldg(t_sw + c), a_sw);
            a_se = fmaf(rv, __ldg(t_se + c), a_se);
          }
        }
        const float part = a_nw * nw + a_ne * ne + a_sw * sw + a_se * se;
        const float tot = warp_sum(part);
        if (lane == j) res = tot;
      }
      s_corr[lane * 33 + pl] = valid ? fdiv(res, scale) : 0.f;
      s_mask[lane * 33 + pl] = valid ? 1.f : 0.f;
    }
    __syncthreads();
    {  // coalesced 128-byte rows: one plane x 32 pixels
      const int col = t & 31;
      for (int rrow = t >> 5; rrow < 32; rrow += kWarps) {
        const int s = sc + rrow;
        if (s < A.s_end && p0 + col < P) {
          const size_t o = (out_base + A.out_s0 + (s - A.s_begin)) * P + p0 + col;
          A.corr[o] = s_corr[rrow * 33 + col];
          A.mask[o] = s_mask[rrow * 33 + col];
        }
      }
    }
    __syncthreads();
  }
}

// ---------------------------------------------------------------------------------------------------------
// WarpOnlyCorr (planesweep_corr.py:107-140): source features warped to every plane, no reduction.
// Thread = one output pixel (n, s, y, x), x fastest -> coalesced stores of warped[n,s,c,y,x] per channel.
// ---------------------------------------------------------------------------------------------------------
struct WarpArgs {
  const float* src;  // [N,C,hs,ws]
  const float* Kk;
  const float* Ks;
  const float* T;
  const float* invd;
  long long ist[4];
  int N, C, h, w, hs, ws, S, s_begin, s_end, norm, out_S, out_s0;
  float* warped;  // [N,out_S,C,h,w]
  float* mask;    // [N,out_S,h,w]
};

__global__ void __launch_bounds__(256) psweep_warp_kernel(const WarpArgs A) {
  const int P = A.h * A.w;
  const int Sl = A.s_end - A.s_begin;
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const int n = blockIdx.y;
  if (idx >= (long long)Sl * P) return;
  const int sl = (int)(idx / P), p = (int)(idx - (long long)sl * P);
  const int s = A.s_begin + sl;
  const int yi = p / A.w, xi = p - yi * A.w;
  const Epi E = epi_from_calib(A.Kk + n * 9, A.Ks + n * 9, A.T + n * 16, A.h, A.w, A.hs, A.ws);
  const EpiPix ep = epi_pixel(E, xi, yi);
  const float d = A.invd[n * A.ist[0] + s * A.ist[1] + yi * A.ist[2] + xi * A.ist[3]];
  float us, vs;
  bool geom;
  epi_sample(E, ep, d, us, vs, geom);
  const int hs = A.hs, ws = A.ws;
  const Bilinear b = bilinear_setup(pix_to_unnorm(us, ws), pix_to_unnorm(vs, hs), hs, ws);
  const bool m = sampling_mask(b.msum);
  const float mf = m ? 1.f : 0.f;
  const int x0 = sat_index(b.x0f, ws), y0 = sat_index(b.y0f, hs);
  const int x0c = clampi(x0, 0, ws - 1), x1c = clampi(x0 + 1, 0, ws - 1);
  const int y0c = clampi(y0, 0, hs - 1), y1c = clampi(y0 + 1, 0, hs - 1);
  const int o_nw = y0c * ws + x0c, o_ne = y0c * ws + x1c, o_sw = y1c * ws + x0c, o_se = y1c * ws + x1c;
  const int Ps = hs * ws;
  const float* src = A.src + (size_t)n * A.C * Ps;
  const size_t so = (size_t)n * A.out_S + A.out_s0 + sl;
  A.mask[so * P + p] = mf;
  float* out = A.warped + so * A.C * P + p;

  float in_nw = 1.f, in_ne = 1.f, in_sw = 1.f, in_se = 1.f;  // divisors of the four corner texels
  if (A.norm == RMVD_NORM_BEFORE) {                          // normalize(feat_src, dim=1), :128-129
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
    for (int c = 0; c < A.C; ++c) {
      const float* sc = src + (size_t)c * Ps;
      const float q0 = __ldg(sc + o_nw), q1 = __ldg(sc + o_ne), q2 = __ldg(sc + o_sw), q3 = __ldg(sc + o_se);
      a0 = fmaf(q0, q0, a0); a1 = fmaf(q1, q1, a1); a2 = fmaf(q2, q2, a2); a3 = fmaf(q3, q3, a3);
    }
    in_nw = sqrtf(a0) + 1e-9f; in_ne = sqrtf(a1) + 1e-9f; in_sw = sqrtf(a2) + 1e-9f; in_se = sqrtf(a3) + 1e-9f;
  }
  const bool nb = (A.norm == RMVD_NORM_BEFORE);
  float onorm = 1.f;
  if (A.norm == RMVD_NORM_AFTER) {  // normalize(warped, dim=2), :135-136
    float acc = 0.f;
    for (int c = 0; c < A.C; ++c) {
      const float* sc = src + (size_t)c * Ps;
      const float val = __ldg(sc + o_nw) * b.nw + __ldg(sc + o_ne) * b.ne + __ldg(sc + o_sw) * b.sw + __ldg(sc + o_se) * b.se;
      acc = fmaf(val, val, acc);
    }
    onorm = sqrtf(acc) + 1e-9f;
  }
  // 8 channels per trip: 32 independent gathers in flight per thread (the loop is latency-bound otherwise), streaming stores
  const float wnw = b.nw * mf, wne = b.ne * mf, wsw = b.sw * mf, wse = b.se * mf;   // mf is 0 or 1: (sum q w) * mf, :138
  const bool plain = !nb && (A.norm != RMVD_NORM_AFTER);
  int c = 0;
  if (plain) {
    for (; c + 8 <= A.C; c += 8) {
      float q[8][4];
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const float* sc = src + (size_t)(c + k) * Ps;
        q[k][0] = __ldg(sc + o_nw); q[k][1] = __ldg(sc + o_ne); q[k][2] = __ldg(sc + o_sw); q[k][3] = __ldg(sc + o_se);
      }
#pragma unroll
      for (int k = 0; k < 8; ++k) __stcs(out + (size_t)(c + k) * P, q[k][0] * wnw + q[k][1] * wne + q[k][2] * wsw + q[k][3] * wse);
    }
  }
  for (; c < A.C; ++c) {
    const float* sc = src + (size_t)c * Ps;
    float q0 = __ldg(sc + o_nw), q1 = __ldg(sc + o_ne), q2 = __ldg(sc + o_sw), q3 = __ldg(sc + o_se);
    if (nb) { q0 = fdiv(q0, in_nw); q1 = fdiv(q1, in_ne); q2 = fdiv(q2, in_sw); q3 = fdiv(q3, in_se); }
    float val = q0 * b.nw + q1 * b.ne + q2 * b.sw + q3 * b.se;
    if (A.norm == RMVD_NORM_AFTER) val = fdiv(val, onorm);
    out[(size_t)c * P] = val * mf;  // warped * warping_mask, :138
  }
}

// ---------------------------------------------------------------------------------------------------------
// warp() grid branch (planesweep_corr.py:49-104) for arbitrary images: thread = one output pixel.
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) warp_grid_kernel(const float* __restrict__ x, const float* __restrict__ grid,
                                                        int C, int hs, int ws, int Po, int border, int x_repeat,
                                                        float* __restrict__ out, float* __restrict__ mask) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  const int k = blockIdx.y;
  if (p >= Po) return;
  const float us = grid[((size_t)k * 2) * Po + p], vs = grid[((size_t)k * 2 + 1) * Po + p];
  float ix = pix_to_unnorm(us, ws), iy = pix_to_unnorm(vs, hs);
  if (border) {  // ATen clip_coordinates: min(size - 1, max(coord, 0)); with the CPU's std::min / std::max a nan becomes size - 1
    ix = isnan(ix) ? (float)(ws - 1) : fminf((float)(ws - 1), fmaxf(ix, 0.f));
    iy = isnan(iy) ? (float)(hs - 1) : fminf((float)(hs - 1), fmaxf(iy, 0.f));
  }
  const Bilinear b = bilinear_setup(ix, iy, hs, ws);
  const int x0 = sat_index(b.x0f, ws), y0 = sat_index(b.y0f, hs);
  const int x0c = clampi(x0, 0, ws - 1), x1c = clampi(x0 + 1, 0, ws - 1);
  const int y0c = clampi(y0, 0, hs - 1), y1c = clampi(y0 + 1, 0, hs - 1);
  const int o_nw = y0c * ws + x0c, o_ne = y0c * ws + x1c, o_sw = y1c * ws + x0c, o_se = y1c * ws + x1c;
  const int Ps = hs * ws;
  const float* xs = x + (size_t)(k / x_repeat) * C * Ps;   // warp_multi: x_repeat consecutive grids sample the same image (:33)
  for (int c = 0; c < C; ++c) {
    const float* sc = xs + (size_t)c * Ps;
    out[((size_t)k * C + c) * Po + p] =
        __ldg(sc + o_nw) * b.nw + __ldg(sc + o_ne) * b.ne + __ldg(sc + o_sw) * b.sw + __ldg(sc + o_se) * b.se;
  }
  mask[(size_t)k * Po + p] = border ? 1.f : (sampling_mask(b.msum) ? 1.f : 0.f);  // :95-102
}

// ---------------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------------
static size_t corr_smem_bytes(int C) { return (size_t)(kTilePx * (C + 4) + 2 * 32 * 33) * sizeof(float); }

template <int kC4, bool kFromGrid>
static int launch_corr(const PsweepArgs& A, int nviews, cudaStream_t st) {
  const size_t smem = corr_smem_bytes(A.C);
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(psweep_corr_kernel<kC4, kFromGrid>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
  }
  const int P = A.h * A.w;
  dim3 grid((P + kTilePx - 1) / kTilePx, A.N, nviews);
  psweep_corr_kernel<kC4, kFromGrid><<<grid, kWarps * 32, smem, st>>>(A);
  return (int)cudaGetLastError();
}

template <bool kFromGrid>
static int dispatch_corr(const PsweepArgs& A, int nviews, cudaStream_t st) {
  switch (A.C) {
    case 128: return launch_corr<1, kFromGrid>(A, nviews, st);
    case 256: return launch_corr<2, kFromGrid>(A, nviews, st);
    case 384: return launch_corr<3, kFromGrid>(A, nviews, st);
    case 512: return launch_corr<4, kFromGrid>(A, nviews, st);
    default: return launch_corr<0, kFromGrid>(A, nviews, st);
  }
}

static int repack(const float* src, float* dst, int N, int C, int P, int l2norm, cudaStream_t st) {
  dim3 grid((P + 31) / 32, N);
  nchw_to_nhwc_kernel<<<grid, 256, 0, st>>>(src, dst, C, P, l2norm);
  return (int)cudaGetLastError();
}

}  // namespace rmvd

using namespace rmvd;

extern "C" size_t rmvd_psweep_workspace_bytes(int N, int C, int h, int w, int V, const int* hs, const int* ws) {
  size_t tot = (psweep_ws_key_bytes(N, C, h, w) + 255) / 256 * 256;          // key copy (whole tiles of 64 pixels)
  for (int v = 0; v < V; ++v) tot += (psweep_ws_src_bytes(N, C, hs[v], ws[v]) + 255) / 256 * 256;   // rows padded to 8 texels
  return tot + 256;                                                                                  // work-item counter
}

static int psweep_corr_fwd_impl(const float* feat_key, const float* const* feat_src, const float* intrinsics_key,
                                const float* const* intrinsics_src, const float* const* src_to_key,
                                const float* invdepths, const int64_t* inv_stride, int N, int C, int h, int w,
                                int V, const int* hs, const int* ws, int S, int s_begin, int s_end,
                                int normalize_mode, float* corr, float* mask, int out_S, int out_s0,
                                int32_t* dbg_ix0, int32_t* dbg_iy0, uint8_t* dbg_geom, void* workspace,
                                size_t workspace_bytes, void* stream, const PsweepPeers* peers, bool act = false, float act_slope = 1.f,
                                float* key_act_out = nullptr) {
  if (!feat_key || !feat_src || !intrinsics_key || !intrinsics_src || !src_to_key || !invdepths || !inv_stride ||
      !hs || !ws || !corr || !mask)
    return RMVD_E_NULL;
  if (N <= 0 || C <= 0 || h <= 0 || w <= 0 || V <= 0 || S <= 0 || out_S < 0) return RMVD_E_SHAPE;
  if (s_begin < 0 || s_end > S || s_begin > s_end || out_s0 < 0 || out_s0 + (s_end - s_begin) > out_S)
    return RMVD_E_RANGE;
  if (out_S == 0 && (s_begin != s_end || peers)) return RMVD_E_SHAPE;   // an empty slab of a single-GPU call is a no-op (below)
  if (normalize_mode != RMVD_NORM_NONE && normalize_mode != RMVD_NORM_DIM && normalize_mode != RMVD_NORM_BEFORE)
    return RMVD_E_MODE;
  if ((dbg_ix0 == nullptr) != (dbg_iy0 == nullptr)) return RMVD_E_NULL;
  for (int v = 0; v < V; ++v) {
    if (!feat_src[v] || !intrinsics_src[v] || !src_to_key[v]) return RMVD_E_NULL;
    if (hs[v] <= 0 || ws[v] <= 0 || hs[v] >= 32768 || ws[v] >= 32768) return RMVD_E_SHAPE;
  }
  if (workspace_bytes < rmvd_psweep_workspace_bytes(N, C, h, w, V, hs, ws) || !workspace) return RMVD_E_WORKSPACE;
  if (s_begin == s_end && !peers) return RMVD_OK;     // (an empty slab of the peer mode still takes part in the handshake)
  cudaStream_t st = (cudaStream_t)stream;

  // kernel choice: persistent tcgen05 kernel (psweep_ws.cu: C % 32 == 0, C <= 256, per-plane hypotheses) > banded fp32
  // kernel (C % 32 == 0, source map <= 65536 texels) > general direct kernel.  RMVD_PSWEEP_KERNEL=ws|band|direct restricts
  // the choice (A/B runs only).
  static const char* const force = getenv("RMVD_PSWEEP_KERNEL");   // read once per process
  const bool allow_ws = !force || strcmp(force, "ws") == 0;
  const bool allow_band = !force || strcmp(force, "band") == 0;
  bool band = (C % 32 == 0) && allow_band;
  for (int v = 0; v < V; ++v) band = band && ((long long)hs[v] * ws[v] <= 65536);
  bool wsk = allow_ws;
  {
    PsweepArgs T{};
    T.C = C; T.h = h; T.w = w;
    for (int k = 0; k < 4; ++k) T.ist[k] = inv_stride[k];
    wsk = wsk && psweep_ws_eligible(T, 0) && (long long)N * (V < kMaxViews ? V : kMaxViews) <= 65535;
  }
  if (peers && (!wsk || V > kMaxViews)) return RMVD_E_MODE;
  if (act && !wsk) return RMVD_E_MODE;                 // the fused activation lives in the split copies of the persistent kernel

  char* wsp = (char*)workspace;
  unsigned* counter = (unsigned*)(wsp + rmvd_psweep_workspace_bytes(N, C, h, w, V, hs, ws) - 256);
  float* key_nhwc = (float*)wsp;
  wsp += (psweep_ws_key_bytes(N, C, h, w) + 255) / 256 * 256;
  const int l2 = normalize_mode == RMVD_NORM_BEFORE;
  if (wsk) {
    int rc = repack_key_tiles(feat_key, key_nhwc, N, C, h, w, l2, st, act, act_slope, key_act_out);
    if (rc) return rc;
  } else if (band) {
    int rc = repack(feat_key, key_nhwc, N, C, h * w, l2, st);
    if (rc) return rc;
  }
  for (int v0 = 0; v0 < V; v0 += kMaxViews) {
    const int nv = (V - v0 < kMaxViews) ? (V - v0) : kMaxViews;
    PsweepArgs A{};
    A.key = feat_key;
    A.Kk = intrinsics_key;
    A.invd = invdepths;
    for (int k = 0; k < 4; ++k) A.ist[k] = inv_stride[k];
    const float* rp_src[1 + kMaxViews];
    void* rp_dst[1 + kMaxViews];
    int rp_n = 0;
    for (int j = 0; j < nv; ++j) {
      const int v = v0 + j;
      float* dst = (float*)wsp;
      wsp += (psweep_ws_src_bytes(N, C, hs[v], ws[v]) + 255) / 256 * 256;
      if (wsk) {
        rp_src[rp_n] = feat_src[v]; rp_dst[rp_n] = dst; ++rp_n;
      } else {
        int rc = repack(feat_src[v], dst, N, C, hs[v] * ws[v], l2, st);
        if (rc) return rc;
      }
      A.src[j] = dst;
      A.Ks[j] = intrinsics_src[v];
      A.T[j] = src_to_key[v];
      A.hs[j] = hs[v];
      A.ws[j] = ws[v];
    }
    if (wsk) {
      int rc = repack_grouped_batch(rp_src, rp_dst, hs + v0, ws + v0, rp_n, N, C, l2, counter, st, act, act_slope);
      if (rc) return rc;
    }
    A.N = N; A.C = C; A.h = h; A.w = w; A.S = S; A.s_begin = s_begin; A.s_end = s_end;
    A.norm = normalize_mode; A.out_S = out_S; A.out_s0 = out_s0; A.v0 = v0;
    A.corr = corr; A.mask = mask; A.dbg_ix0 = dbg_ix0; A.dbg_iy0 = dbg_iy0; A.dbg_geom = dbg_geom;
    int rc = wsk ? launch_psweep_ws(A, key_nhwc, nv, counter, peers, st)
                : band ? launch_psweep_band(A, key_nhwc, nv, st) : dispatch_corr<false>(A, nv, st);
    if (rc) return rc;
  }
  return RMVD_OK;
}

extern "C" int rmvd_psweep_corr_fwd(const float* feat_key, const float* const* feat_src, const float* intrinsics_key,
                                    const float* const* intrinsics_src, const float* const* src_to_key,
                                    const float* invdepths, const int64_t* inv_stride, int N, int C, int h, int w,
                                    int V, const int* hs, const int* ws, int S, int s_begin, int s_end,
                                    int normalize_mode, float* corr, float* mask, int out_S, int out_s0,
                                    int32_t* dbg_ix0, int32_t* dbg_iy0, uint8_t* dbg_geom, void* workspace,
                                    size_t workspace_bytes, void* stream) {
  return psweep_corr_fwd_impl(feat_key, feat_src, intrinsics_key, intrinsics_src, src_to_key, invdepths, inv_stride, N, C, h, w, V, hs, ws,
                              S, s_begin, s_end, normalize_mode, corr, mask, out_S, out_s0, dbg_ix0, dbg_iy0, dbg_geom, workspace,
                              workspace_bytes, stream, nullptr);
}

extern "C" int rmvd_psweep_corr_fwd_act(const float* pre_key, const float* const* pre_src, float act_slope, float* key_act_out,
                                        const float* intrinsics_key, const float* const* intrinsics_src, const float* const* src_to_key,
                                        const float* invdepths, const int64_t* inv_stride, int N, int C, int h, int w,
                                        int V, const int* hs, const int* ws, int S, int s_begin, int s_end,
                                        int normalize_mode, float* corr, float* mask, int out_S, int out_s0, void* workspace,
                                        size_t workspace_bytes, void* stream) {
  if (!(act_slope == act_slope)) return RMVD_E_MODE;
  return psweep_corr_fwd_impl(pre_key, pre_src, intrinsics_key, intrinsics_src, src_to_key, invdepths, inv_stride, N, C, h, w, V, hs, ws,
                              S, s_begin, s_end, normalize_mode, corr, mask, out_S, out_s0, nullptr, nullptr, nullptr, workspace,
                              workspace_bytes, stream, nullptr, true, act_slope, key_act_out);
}

extern "C" int rmvd_psweep_corr_fwd_peers(const float* feat_key, const float* const* feat_src, const float* intrinsics_key,
                                          const float* const* intrinsics_src, const float* const* src_to_key,
                                          const float* invdepths, const int64_t* inv_stride, int N, int C, int h, int w,
                                          int V, const int* hs, const int* ws, int S, int s_begin, int s_end,
                                          int normalize_mode, int out_S, int out_s0, int num_peers, int rank,
                                          float* const* corr_peers, float* const* mask_peers, uint32_t* const* flag_peers,
                                          uint32_t epoch, int partition, float* corr_multicast, float* mask_multicast, void* workspace,
                                          size_t workspace_bytes, void* stream) {
  if (!corr_peers || !mask_peers || !flag_peers) return RMVD_E_NULL;
  if (num_peers < 1 || num_peers > RMVD_MAX_PEERS || rank < 0 || rank >= num_peers) return RMVD_E_RANGE;
  if (partition != RMVD_PARTITION_PLANES && partition != RMVD_PARTITION_TILES) return RMVD_E_MODE;
  PsweepPeers P{};
  if ((corr_multicast == nullptr) != (mask_multicast == nullptr)) return RMVD_E_NULL;
  P.n = num_peers; P.rank = rank; P.epoch = epoch; P.by_tiles = (partition == RMVD_PARTITION_TILES);
  P.corr_mc = corr_multicast; P.mask_mc = mask_multicast;
  for (int g = 0; g < num_peers; ++g) {
    if (!corr_peers[g] || !mask_peers[g] || !flag_peers[g]) return RMVD_E_NULL;
    P.corr[g] = corr_peers[g]; P.mask[g] = mask_peers[g]; P.flags[g] = flag_peers[g];
  }
  return psweep_corr_fwd_impl(feat_key, feat_src, intrinsics_key, intrinsics_src, src_to_key, invdepths, inv_stride, N, C, h, w, V, hs, ws,
                              S, s_begin, s_end, normalize_mode, P.corr[rank], P.mask[rank], out_S, out_s0, nullptr, nullptr, nullptr,
                              workspace, workspace_bytes, stream, &P);
}

// Push transport of the depth-slab gather (psweep_push.cu): the sweep stores locally and publishes finished chunks, a pusher kernel on
// `push_stream` forwards them to the peers from the SMs the sweep leaves free and ends with the ranks' flag handshake.
static unsigned push_queue_cap(int N, int h, int w, int V, int planes) {
  int tw, th;
  psweep_ws_tile_shape(h, w, &tw, &th);
  const long long tiles = (long long)((w + tw - 1) / tw) * ((h + th - 1) / th);
  const long long cap = tiles * N * V * ((planes + 3) / 4 + 1);      // a chunk holds at least 4 planes (or the item's last ones)
  return (unsigned)(((cap > 0x3fffffffLL ? 0x3fffffffLL : cap) + 3) & ~3LL);
}

extern "C" size_t rmvd_psweep_push_queue_bytes(int N, int h, int w, int V, int planes) {
  if (N <= 0 || h <= 0 || w <= 0 || V <= 0 || planes <= 0) return 0;
  return 16 + (size_t)push_queue_cap(N, h, w, V, planes) * 20;
}

extern "C" int rmvd_psweep_corr_fwd_push(const float* feat_key, const float* const* feat_src, const float* intrinsics_key,
                                         const float* const* intrinsics_src, const float* const* src_to_key, const float* invdepths,
                                         const int64_t* inv_stride, int N, int C, int h, int w, int V, const int* hs, const int* ws, int S,
                                         int s_begin, int s_end, int normalize_mode, int out_S, int out_s0, int num_peers, int rank,
                                         float* const* corr_peers, float* const* mask_peers, uint32_t* const* flag_peers, uint32_t epoch,
                                         void* queue, size_t queue_bytes, int push_ctas, void* workspace, size_t workspace_bytes,
                                         void* stream, void* push_stream) {
  if (!corr_peers || !mask_peers || !flag_peers || !queue) return RMVD_E_NULL;
  if (num_peers < 1 || num_peers > RMVD_MAX_PEERS || rank < 0 || rank >= num_peers || push_ctas < 1 || push_ctas > 64) return RMVD_E_RANGE;
  if (s_begin < 0 || s_end > S || s_begin > s_end) return RMVD_E_RANGE;
  if (stream == push_stream) return RMVD_E_MODE;            // the pusher waits for the sweep: they must be able to run side by side
  const size_t need = rmvd_psweep_push_queue_bytes(N, h, w, V, s_end > s_begin ? s_end - s_begin : 1);
  if (queue_bytes < need || (uintptr_t)queue % 16 != 0) return RMVD_E_WORKSPACE;
  PsweepPeers P{};
  P.n = num_peers; P.rank = rank; P.epoch = epoch;
  for (int g = 0; g < num_peers; ++g) {
    if (!corr_peers[g] || !mask_peers[g] || !flag_peers[g]) return RMVD_E_NULL;
    P.corr[g] = corr_peers[g]; P.mask[g] = mask_peers[g]; P.flags[g] = flag_peers[g];
  }
  P.q_head = (unsigned*)queue;
  P.q_cap = (unsigned)((queue_bytes - 16) / 20) & ~3u;
  P.push_ctas = push_ctas;
  // the sweep first: only a sweep that was launched will ever close the queue (a pusher without one would wait for its time-out)
  int rc = psweep_corr_fwd_impl(feat_key, feat_src, intrinsics_key, intrinsics_src, src_to_key, invdepths, inv_stride, N, C, h, w, V, hs, ws,
                                S, s_begin, s_end, normalize_mode, P.corr[rank], P.mask[rank], out_S, out_s0, nullptr, nullptr, nullptr,
                                workspace, workspace_bytes, stream, &P);
  if (rc) return rc;
  return launch_psweep_push(P, h, w, out_S, (cudaStream_t)push_stream);
}

// Copy-engine transport of the depth-slab gather: `rows` runs of `row_bytes` bytes, `pitch_bytes` apart in both buffers (a rank's
// planes [lo, hi) of several [S,h,w] volumes), device to (peer) device.  One cudaMemcpy2DAsync: the copy engines move it at NVLink
// speed without occupying an SM, so it overlaps the persistent plane-sweep kernel of the next view group (parallel.SlabGather).
extern "C" int rmvd_copy_rows(void* dst, const void* src, size_t pitch_bytes, size_t row_bytes, int rows, void* stream) {
  if (!dst || !src) return RMVD_E_NULL;
  if (rows < 0 || row_bytes > pitch_bytes) return RMVD_E_SHAPE;
  if (rows == 0 || row_bytes == 0) return RMVD_OK;
  return (int)cudaMemcpy2DAsync(dst, pitch_bytes, src, pitch_bytes, row_bytes, (size_t)rows, cudaMemcpyDeviceToDevice, (cudaStream_t)stream);
}

// ---------------------------------------------------------------------------------------------------------
// rmvd_psweep_corr_fwd_static: the forward call for callers that may not allocate (CUDA graph capture, serving loops,
// harnesses that forbid cudaMalloc).  The scratch of rmvd_psweep_corr_fwd -- the repacked key / source maps and the
// work-item counter -- lives in this file-scope device array instead of a caller's workspace: RMVD_STATIC_SCRATCH_BYTES
// of every GPU that loads the library, held until the process exits.  The array's name is a host-side stand-in in host
// code, not a device address; the address is taken once per device with cudaGetSymbolAddress (that device current) and
// cached.  Nothing in the scratch has to start at zero except the work-item counters, which the repack launch of every
// call zeroes; everything else is written before it is read.  All calls share the array: calls that overlap on the GPU
// (two streams, two threads) race.
static __device__ __align__(256) unsigned char g_static_scratch[RMVD_STATIC_SCRATCH_BYTES];

static int static_scratch_address(void** out) {
  constexpr int kMaxDevices = 64;
  static std::mutex mu;
  static void* addr[kMaxDevices] = {};
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  if (dev < 0 || dev >= kMaxDevices) return RMVD_E_RANGE;
  std::lock_guard<std::mutex> lock(mu);
  if (!addr[dev]) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_static_scratch);
    if (e != cudaSuccess) return (int)e;
    if (!p) return (int)cudaErrorInvalidSymbol;
    addr[dev] = p;
  }
  *out = addr[dev];
  return RMVD_OK;
}

extern "C" int rmvd_psweep_corr_fwd_static(const float* feat_key, const float* const* feat_src, const float* intrinsics_key,
                                           const float* const* intrinsics_src, const float* const* src_to_key,
                                           const float* invdepths, const int64_t* inv_stride, int N, int C, int h, int w,
                                           int V, const int* hs, const int* ws, int S, int s_begin, int s_end,
                                           int normalize_mode, float* corr, float* mask, int out_S, int out_s0, void* stream) {
  if (!hs || !ws) return RMVD_E_NULL;
  if (N <= 0 || C <= 0 || h <= 0 || w <= 0 || V <= 0) return RMVD_E_SHAPE;
  for (int v = 0; v < V; ++v)
    if (hs[v] <= 0 || ws[v] <= 0 || hs[v] >= 32768 || ws[v] >= 32768) return RMVD_E_SHAPE;
  const size_t need = rmvd_psweep_workspace_bytes(N, C, h, w, V, hs, ws);
  if (need > (size_t)RMVD_STATIC_SCRATCH_BYTES) return RMVD_E_STATIC_LIMIT;
  void* scratch = nullptr;
  int rc = static_scratch_address(&scratch);
  if (rc) return rc;
  return psweep_corr_fwd_impl(feat_key, feat_src, intrinsics_key, intrinsics_src, src_to_key, invdepths, inv_stride, N, C, h, w, V, hs, ws,
                              S, s_begin, s_end, normalize_mode, corr, mask, out_S, out_s0, nullptr, nullptr, nullptr, scratch,
                              (size_t)RMVD_STATIC_SCRATCH_BYTES, stream, nullptr);
}

extern "C" int rmvd_corr_grid_fwd(const float* feat_key, const float* feat_src, const float* grids,
                                  const uint8_t* geom_mask, int N, int C, int h, int w, int hs, int ws, int S,
                                  int normalize_mode, int padding_border, float* corr, float* mask, void* workspace,
                                  size_t workspace_bytes, void* stream) {
  if (!feat_key || !feat_src || !grids || !corr || !mask) return RMVD_E_NULL;
  if (N <= 0 || C <= 0 || h <= 0 || w <= 0 || S <= 0 || hs <= 0 || ws <= 0 || hs >= 32768 || ws >= 32768)
    return RMVD_E_SHAPE;
  if (normalize_mode != RMVD_NORM_NONE && normalize_mode != RMVD_NORM_DIM && normalize_mode != RMVD_NORM_BEFORE)
    return RMVD_E_MODE;
  if (!workspace || workspace_bytes < rmvd_psweep_workspace_bytes(N, C, h, w, 1, &hs, &ws)) return RMVD_E_WORKSPACE;
  cudaStream_t st = (cudaStream_t)stream;
  int rc = repack(feat_src, (float*)workspace, N, C, hs * ws, normalize_mode == RMVD_NORM_BEFORE, st);
  if (rc) return rc;
  PsweepArgs A{};
  A.key = feat_key;
  A.src[0] = (const float*)workspace;
  A.grids = grids;
  A.grid_mask = geom_mask;
  A.border = padding_border;
  A.hs[0] = hs; A.ws[0] = ws;
  A.N = N; A.C = C; A.h = h; A.w = w; A.S = S; A.s_begin = 0; A.s_end = S;
  A.norm = normalize_mode; A.out_S = S; A.out_s0 = 0; A.v0 = 0;
  A.corr = corr; A.mask = mask;
  return dispatch_corr<true>(A, 1, st);
}

// ---------------------------------------------------------------------------------------------------------
// What the reference block leaves on itself after forward() (planesweep_corr.py:389-394, :436-462, :489-512): the homogeneous
// epipolar terms of every key pixel and the sampling points us / vs / zs / geometric mask of every (plane, pixel), materialised on
// request only (the fused correlation never writes them).  Thread = (plane, pixel); every value through the exact chain of common.cuh.
// ---------------------------------------------------------------------------------------------------------
namespace rmvd {
__global__ void __launch_bounds__(256) sampling_points_kernel(const float* __restrict__ Kk, const float* __restrict__ Ks, const float* __restrict__ T,
                                                              const float* __restrict__ invd, long long i0, long long i1, long long i2, long long i3,
                                                              int h, int w, int hs, int ws, int S, float* __restrict__ us, float* __restrict__ vs,
                                                              float* __restrict__ zs, uint8_t* __restrict__ mask, float* __restrict__ pix,
                                                              float* __restrict__ mvec) {
  const int n = blockIdx.y, P = h * w;
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (long long)S * P) return;
  const int s = (int)(idx / P), p = (int)(idx - (long long)s * P);
  const int yi = p / w, xi = p - yi * w;
  const Epi E = epi_from_calib(Kk + n * 9, Ks + n * 9, T + n * 16, h, w, hs, ws);
  const EpiPix px = epi_pixel(E, xi, yi);
  const float d = invd[n * i0 + s * i1 + yi * i2 + xi * i3];
  float u, v;
  bool geom;
  epi_sample(E, px, d, u, v, geom);
  const size_t o = ((size_t)n * S + s) * P + p;
  if (us) { us[o] = u; vs[o] = v; zs[o] = fdiv(1.0f, d); mask[o] = geom ? 1 : 0; }
  if (pix && s == 0) {
    pix[((size_t)n * 3 + 0) * P + p] = px.u_h; pix[((size_t)n * 3 + 1) * P + p] = px.v_h; pix[((size_t)n * 3 + 2) * P + p] = px.k_h;
    if (p == 0) { mvec[n * 3 + 0] = E.e; mvec[n * 3 + 1] = E.i; mvec[n * 3 + 2] = E.m; }
  }
}
}  // namespace rmvd

extern "C" int rmvd_sampling_points_fwd(const float* intrinsics_key, const float* intrinsics_src, const float* src_to_key, const float* invdepths,
                                        const int64_t* inv_stride, int N, int h, int w, int hs, int ws, int S, float* us, float* vs, float* zs,
                                        uint8_t* geom_mask, float* coeff_pix, float* coeff_m, void* stream) {
  if (!intrinsics_key || !intrinsics_src || !src_to_key || !invdepths || !inv_stride) return RMVD_E_NULL;
  if ((us == nullptr) != (vs == nullptr) || (us == nullptr) != (zs == nullptr) || (us == nullptr) != (geom_mask == nullptr)) return RMVD_E_NULL;
  if ((coeff_pix == nullptr) != (coeff_m == nullptr) || (!us && !coeff_pix)) return RMVD_E_NULL;
  if (N <= 0 || h <= 0 || w <= 0 || hs <= 0 || ws <= 0 || S <= 0 || N > 65535) return RMVD_E_SHAPE;
  const long long tot = (long long)(us ? S : 1) * h * w;
  dim3 grid((unsigned)((tot + 255) / 256), N);
  rmvd::sampling_points_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(intrinsics_key, intrinsics_src, src_to_key, invdepths, inv_stride[0], inv_stride[1],
                                                                      inv_stride[2], inv_stride[3], h, w, hs, ws, us ? S : 1, us, vs, zs, geom_mask,
                                                                      coeff_pix, coeff_m);
  return (int)cudaGetLastError();
}

extern "C" int rmvd_psweep_warp_fwd(const float* feat_src, const float* intrinsics_key, const float* intrinsics_src,
                                    const float* src_to_key, const float* invdepths, const int64_t* inv_stride,
                                    int N, int C, int h, int w, int hs, int ws, int S, int s_begin, int s_end,
                                    int normalize_mode, float* warped, float* mask, int out_S, int out_s0,
                                    void* stream) {
  if (!feat_src || !intrinsics_key || !intrinsics_src || !src_to_key || !invdepths || !inv_stride || !warped || !mask)
    return RMVD_E_NULL;
  if (N <= 0 || C <= 0 || h <= 0 || w <= 0 || S <= 0 || hs <= 0 || ws <= 0 || out_S <= 0) return RMVD_E_SHAPE;
  if (s_begin < 0 || s_end > S || s_begin > s_end || out_s0 < 0 || out_s0 + (s_end - s_begin) > out_S)
    return RMVD_E_RANGE;
  if (normalize_mode < RMVD_NORM_NONE || normalize_mode > RMVD_NORM_AFTER) return RMVD_E_MODE;
  if (s_begin == s_end) return RMVD_OK;
  WarpArgs A{};
  A.src = feat_src; A.Kk = intrinsics_key; A.Ks = intrinsics_src; A.T = src_to_key; A.invd = invdepths;
  for (int k = 0; k < 4; ++k) A.ist[k] = inv_stride[k];
  A.N = N; A.C = C; A.h = h; A.w = w; A.hs = hs; A.ws = ws; A.S = S; A.s_begin = s_begin; A.s_end = s_end;
  A.norm = normalize_mode; A.out_S = out_S; A.out_s0 = out_s0; A.warped = warped; A.mask = mask;
  const long long tot = (long long)(s_end - s_begin) * h * w;
  dim3 grid((unsigned)((tot + 255) / 256), N);
  psweep_warp_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(A);
  return (int)cudaGetLastError();
}

// Backward of psweep_warp_kernel for normalize False / "dim" (training direction of WarpOnlyCorr, planesweep_corr.py:107-140: autograd
// through grid_sample and `* warping_mask`): thread = (plane, pixel), same locate as the forward; per channel the upstream gradient
// times the 0 / 1 mask goes to the four corner texels with their bilinear weights (zeros padding: corners outside weigh 0) by float
// atomics -- adjacent pixels hit adjacent texels.  Sampling positions carry no gradient.
struct WarpBwdArgs {
  const float* Kk; const float* Ks; const float* T; const float* invd;
  long long ist[4];
  int N, C, h, w, hs, ws, S, s_begin, s_end, out_S, out_s0;
  const float* grad_warped;   // [N,out_S,C,h,w]
  float* grad_src;            // [N,C,hs,ws], zero on entry
};

__global__ void __launch_bounds__(256) psweep_warp_bwd_kernel(const WarpBwdArgs A) {
  const int P = A.h * A.w;
  const int Sl = A.s_end - A.s_begin;
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const int n = blockIdx.y;
  if (idx >= (long long)Sl * P) return;
  const int sl = (int)(idx / P), p = (int)(idx - (long long)sl * P);
  const int s = A.s_begin + sl;
  const int yi = p / A.w, xi = p - yi * A.w;
  const Epi E = epi_from_calib(A.Kk + n * 9, A.Ks + n * 9, A.T + n * 16, A.h, A.w, A.hs, A.ws);
  const EpiPix ep = epi_pixel(E, xi, yi);
  const float d = A.invd[n * A.ist[0] + s * A.ist[1] + yi * A.ist[2] + xi * A.ist[3]];
  float us, vs;
  bool geom;
  epi_sample(E, ep, d, us, vs, geom);
  const int hs = A.hs, ws = A.ws;
  const Bilinear b = bilinear_setup(pix_to_unnorm(us, ws), pix_to_unnorm(vs, hs), hs, ws);
  if (!sampling_mask(b.msum)) return;                    // warped * warping_mask: no gradient through masked samples
  const int x0 = sat_index(b.x0f, ws), y0 = sat_index(b.y0f, hs);
  const int x0c = clampi(x0, 0, ws - 1), x1c = clampi(x0 + 1, 0, ws - 1);
  const int y0c = clampi(y0, 0, hs - 1), y1c = clampi(y0 + 1, 0, hs - 1);
  const int o[4] = {y0c * ws + x0c, y0c * ws + x1c, y1c * ws + x0c, y1c * ws + x1c};
  const float wq[4] = {b.nw, b.ne, b.sw, b.se};
  const int Ps = hs * ws;
  const size_t so = (size_t)n * A.out_S + A.out_s0 + sl;
  const float* g = A.grad_warped + so * A.C * P + p;
  float* gs = A.grad_src + (size_t)n * A.C * Ps;
  for (int c = 0; c < A.C; ++c) {
    const float gv = __ldcs(g + (size_t)c * P);
    if (gv == 0.f) continue;
#pragma unroll
    for (int q = 0; q < 4; ++q)
      if (wq[q] != 0.f) atomicAdd(gs + (size_t)c * Ps + o[q], gv * wq[q]);
  }
}

extern "C" int rmvd_psweep_warp_bwd(const float* intrinsics_key, const float* intrinsics_src, const float* src_to_key,
                                    const float* invdepths, const int64_t* inv_stride, int N, int C, int h, int w, int hs, int ws,
                                    int S, int s_begin, int s_end, int normalize_mode, const float* grad_warped, int out_S, int out_s0,
                                    float* grad_src, void* stream) {
  if (!intrinsics_key || !intrinsics_src || !src_to_key || !invdepths || !inv_stride || !grad_warped || !grad_src) return RMVD_E_NULL;
  if (N <= 0 || C <= 0 || h <= 0 || w <= 0 || S <= 0 || hs <= 0 || ws <= 0 || out_S <= 0 || N > 65535) return RMVD_E_SHAPE;
  if (s_begin < 0 || s_end > S || s_begin > s_end || out_s0 < 0 || out_s0 + (s_end - s_begin) > out_S) return RMVD_E_RANGE;
  if (normalize_mode != RMVD_NORM_NONE && normalize_mode != RMVD_NORM_DIM) return RMVD_E_MODE;   // L2-normalised variants: the caller differentiates
  const cudaError_t e = cudaMemsetAsync(grad_src, 0, (size_t)N * C * hs * ws * sizeof(float), (cudaStream_t)stream);
  if (e != cudaSuccess) return (int)e;
  if (s_begin == s_end) return RMVD_OK;
  WarpBwdArgs A{};
  A.Kk = intrinsics_key; A.Ks = intrinsics_src; A.T = src_to_key; A.invd = invdepths;
  for (int k = 0; k < 4; ++k) A.ist[k] = inv_stride[k];
  A.N = N; A.C = C; A.h = h; A.w = w; A.hs = hs; A.ws = ws; A.S = S; A.s_begin = s_begin; A.s_end = s_end;
  A.out_S = out_S; A.out_s0 = out_s0; A.grad_warped = grad_warped; A.grad_src = grad_src;
  const long long tot = (long long)(s_end - s_begin) * h * w;
  dim3 grid((unsigned)((tot + 255) / 256), N);
  psweep_warp_bwd_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(A);
  return (int)cudaGetLastError();
}

extern "C" int rmvd_warp_grid_fwd(const float* x, const float* grid, int K, int C, int hs, int ws, int ho, int wo,
                                  int padding_border, int x_repeat, float* out, float* mask, void* stream) {
  if (!x || !grid || !out || !mask) return RMVD_E_NULL;
  if (K <= 0 || C <= 0 || hs <= 0 || ws <= 0 || ho <= 0 || wo <= 0 || x_repeat <= 0) return RMVD_E_SHAPE;
  if (padding_border != 0 && padding_border != 1) return RMVD_E_MODE;
  if (K > 65535) return RMVD_E_SHAPE;
  const int Po = ho * wo;
  dim3 g((Po + 255) / 256, K);
  warp_grid_kernel<<<g, 256, 0, (cudaStream_t)stream>>>(x, grid, C, hs, ws, Po, padding_border, x_repeat, out, mask);
  return (int)cudaGetLastError();
}
