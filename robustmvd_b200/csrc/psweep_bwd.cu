// Backward of the robust_mvd plane-sweep correlation for sm_100a.
//
// The reference gets it from autograd through TorchCorr.forward (planesweep_corr.py:152-195): matmul -> /sqrt(C) ->
// grid_sample -> * corr_mask -> * geometric mask; the sampling points carry no gradient (init_coeffs :436,
// get_sampling_invdepths :464 and get_plane_sweep_sampling_points :489 are no_grad).  With
//   corr[n,s,p] = valid * (1/scale) * sum_c key[n,c,p] * sum_k w_k src[n,c,t_k]          (k = 4 bilinear corners)
// the gradients are
//   d key[n,c,p]  = sum_v sum_s g' * sum_k w_k src_v[n,c,t_k]                            (gather)
//   d src_v[n,c,t] = sum_{p,s,k : t_k = t} g' * w_k * key[n,c,p]                         (scatter-add)
// with g' = valid * grad_corr / scale.  valid, t_k and w_k come from the same exactly rounded chain as the forward
// kernels (common.cuh), so a sample contributes to the gradient iff the forward mask is 1.
//
// Data layout: source maps are repacked to channels-last [N, hs*ws, C] (one texel = one contiguous C-vector), the
// gradients are accumulated channels-last in the caller's workspace with vector reductions (red.global.add.v4.f32,
// one 16-byte reduction per lane) and transposed back to the reference's NCHW at the end.
//
// CTA = 32 consecutive key pixels of one (view, sample); warp = 4 of those pixels, lanes split the channel axis.
// Consecutive planes of a pixel that fall into the same source cell are merged before touching memory: the four
// coefficients sum_s g' w_k are accumulated in registers and one gather + one scatter is issued per run of planes.
#include <cstdlib>
#include <cstring>

#include "psweep.cuh"

namespace rmvd {

struct PsweepBwdArgs {
  const float* key;             // [N,C,h,w]
  const float* src[kMaxViews];  // channels-last copies [N,hs*ws,C]
  float* gsrc[kMaxViews];       // channels-last accumulators [N,hs*ws,C] (zeroed) or nullptr
  float* gkey;                  // channels-last accumulator [N,h*w,C] (zeroed) or nullptr
  const float* gcorr;           // [V,N,out_S,h,w]
  const float* Kk;
  const float* Ks[kMaxViews];
  const float* T[kMaxViews];
  const float* invd;
  long long ist[4];
  int hs[kMaxViews], ws[kMaxViews];
  int N, C, h, w, S, s_begin, s_end, norm, out_S, out_s0, v0;
  int planes0, rmax;            // tiled kernel: planes per chunk to start from, texels of a band box
  int zsplit;                   // tiled kernel: CTAs per tile, each sweeping 1/zsplit of the planes (fills the last wave)
};

__device__ __forceinline__ void red_add4(float* p, const float4 v) {
  atomicAdd(reinterpret_cast<float4*>(p), v);  // result unused: RED.E.ADD.F32x4 (sm_90+)
}
__device__ __forceinline__ float4 scale4(const float4 a, float s) { return make_float4(a.x * s, a.y * s, a.z * s, a.w * s); }
__device__ __forceinline__ void fma4(float4& acc, const float4 a, float s) {
  acc.x = fmaf(a.x, s, acc.x);
  acc.y = fmaf(a.y, s, acc.y);
  acc.z = fmaf(a.z, s, acc.z);
  acc.w = fmaf(a.w, s, acc.w);
}

//   kC4 > 0 : C == 128 * kC4, lane owns channels [k*128 + 4*lane, +4) of every 128-channel group;
//   kC4 == 0: any C, scalar channel loop with the key gradient of the tile accumulated in shared memory.
template <int kC4>
__global__ void __launch_bounds__(kWarps * 32) psweep_corr_bwd_kernel(const PsweepBwdArgs A) {
  extern __shared__ __align__(16) float smem[];
  const int C = A.C, Cp = C + 4;
  float* s_ref = smem;                  // [32][Cp] key tile, [pixel][channel]
  float* s_gk = s_ref + kTilePx * Cp;   // [32][Cp] key-gradient tile (kC4 == 0 only)

  const int v = blockIdx.z, n = blockIdx.y;
  const int P = A.h * A.w;
  const int p0 = blockIdx.x * kTilePx;
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int hs = A.hs[v], ws = A.ws[v];
  const float* __restrict__ src = A.src[v] + (size_t)n * hs * ws * C;
  float* __restrict__ gsrc = A.gsrc[v] ? A.gsrc[v] + (size_t)n * hs * ws * C : nullptr;
  float* __restrict__ gkey = A.gkey ? A.gkey + (size_t)n * P * C : nullptr;

  {
    const float* key = A.key + (size_t)n * C * P;
    const int px = t & 31;
    for (int c = t >> 5; c < C; c += kWarps) {
      s_ref[px * Cp + c] = (p0 + px < P) ? key[(size_t)c * P + p0 + px] : 0.f;
      if (kC4 == 0) s_gk[px * Cp + c] = 0.f;
    }
  }
  __syncthreads();

  const Epi E = epi_from_calib(A.Kk + n * 9, A.Ks[v] + n * 9, A.T[v] + n * 16, A.h, A.w, hs, ws);
  const float scale = (A.norm == RMVD_NORM_DIM) ? sqrtf((float)C) : 1.0f;  // corr / sqrt(C), :178-179
  const size_t g_base = ((size_t)(A.v0 + v) * A.N + n) * A.out_S;

  for (int q = 0; q < kPxPerWarp; ++q) {
    const int pl = warp * kPxPerWarp + q;
    const int p = p0 + pl;
    if (p >= P) break;  // warp-uniform
    const int yi = p / A.w, xi = p - yi * A.w;
    const EpiPix ep = epi_pixel(E, xi, yi);
    const float* rr = s_ref + pl * Cp;

    float4 gk[kC4 > 0 ? kC4 : 1];
#pragma unroll
    for (int k = 0; k < (kC4 > 0 ? kC4 : 1); ++k) gk[k] = make_float4(0.f, 0.f, 0.f, 0.f);

    bool have = false;   // a run of planes in one source cell is open
    int cur = 0;         // its packed corner
    float c_nw = 0.f, c_ne = 0.f, c_sw = 0.f, c_se = 0.f;  // sum over the run of g' * w_k

    auto flush = [&]() {
      const int x0c = cur & 0x7fff, y0c = (cur >> 15) & 0x7fff;
      const int dx = (cur >> 30) & 1, dy = (cur >> 31) & 1;
      const size_t o_nw = ((size_t)y0c * ws + x0c) * C;
      const size_t o_ne = o_nw + (size_t)dx * C;
      const size_t o_sw = o_nw + (size_t)dy * ws * C;
      const size_t o_se = o_sw + (size_t)dx * C;
      if (kC4 > 0) {
#pragma unroll
        for (int k = 0; k < (kC4 > 0 ? kC4 : 1); ++k) {
          const int o = k * 128 + lane * 4;
          if (gkey) {
            fma4(gk[k], ldg4(src + o_nw + o), c_nw);
            fma4(gk[k], ldg4(src + o_ne + o), c_ne);
            fma4(gk[k], ldg4(src + o_sw + o), c_sw);
            fma4(gk[k], ldg4(src + o_se + o), c_se);
          }
          if (gsrc) {
            const float4 r = *reinterpret_cast<const float4*>(rr + o);
            if (c_nw != 0.f) red_add4(gsrc + o_nw + o, scale4(r, c_nw));
            if (c_ne != 0.f) red_add4(gsrc + o_ne + o, scale4(r, c_ne));
            if (c_sw != 0.f) red_add4(gsrc + o_sw + o, scale4(r, c_sw));
            if (c_se != 0.f) red_add4(gsrc + o_se + o, scale4(r, c_se));
          }
        }
      } else {
        float* gg = s_gk + pl * Cp;
        for (int c = lane; c < C; c += 32) {
          if (gkey) {
            float a = gg[c];
            a = fmaf(__ldg(src + o_nw + c), c_nw, a);
            a = fmaf(__ldg(src + o_ne + c), c_ne, a);
            a = fmaf(__ldg(src + o_sw + c), c_sw, a);
            a = fmaf(__ldg(src + o_se + c), c_se, a);
            gg[c] = a;
          }
          if (gsrc) {
            const float rv = rr[c];
            if (c_nw != 0.f) atomicAdd(gsrc + o_nw + c, rv * c_nw);
            if (c_ne != 0.f) atomicAdd(gsrc + o_ne + c, rv * c_ne);
            if (c_sw != 0.f) atomicAdd(gsrc + o_sw + c, rv * c_sw);
            if (c_se != 0.f) atomicAdd(gsrc + o_se + c, rv * c_se);
          }
        }
      }
    };

    for (int sc = A.s_begin; sc < A.s_end; sc += 32) {
      const int s = sc + lane;
      bool valid = false;
      int pack = 0;
      float wnw = 0.f, wne = 0.f, wsw = 0.f, wse = 0.f;
      if (s < A.s_end) {
        const float d = A.invd[n * A.ist[0] + s * A.ist[1] + yi * A.ist[2] + xi * A.ist[3]];
        float us, vs;
        bool geom;
        epi_sample(E, ep, d, us, vs, geom);
        const Bilinear b = bilinear_setup(pix_to_unnorm(us, ws), pix_to_unnorm(vs, hs), hs, ws);
        valid = sampling_mask(b.msum) && geom;
        if (valid) {
          const float g = fdiv(A.gcorr[(g_base + A.out_s0 + (s - A.s_begin)) * P + p], scale);
          wnw = b.nw * g; wne = b.ne * g; wsw = b.sw * g; wse = b.se * g;
          const int x0 = sat_index(b.x0f, ws), y0 = sat_index(b.y0f, hs);
          const int x0c = clampi(x0, 0, ws - 1), x1c = clampi(x0 + 1, 0, ws - 1);
          const int y0c = clampi(y0, 0, hs - 1), y1c = clampi(y0 + 1, 0, hs - 1);
          pack = x0c | (y0c << 15) | ((x1c - x0c) << 30) | ((y1c - y0c) << 31);
        }
      }
      unsigned vm = __ballot_sync(kFullMask, valid);
      while (vm) {
        const int j = __ffs(vm) - 1;
        vm &= vm - 1;
        const int pk = __shfl_sync(kFullMask, pack, j);
        const float nw = __shfl_sync(kFullMask, wnw, j), ne = __shfl_sync(kFullMask, wne, j);
        const float sw = __shfl_sync(kFullMask, wsw, j), se = __shfl_sync(kFullMask, wse, j);
        if (have && pk != cur) {
          flush();
          have = false;
        }
        if (!have) {
          have = true;
          cur = pk;
          c_nw = c_ne = c_sw = c_se = 0.f;
        }
        c_nw += nw; c_ne += ne; c_sw += sw; c_se += se;
      }
    }
    if (have) flush();

    if (gkey) {
      float* gp = gkey + (size_t)p * C;
      if (kC4 > 0) {
#pragma unroll
        for (int k = 0; k < (kC4 > 0 ? kC4 : 1); ++k) red_add4(gp + k * 128 + lane * 4, gk[k]);
      } else {
        __syncwarp();
        const float* gg = s_gk + pl * Cp;
        for (int c = lane; c < C; c += 32) atomicAdd(gp + c, gg[c]);
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------------------
// Tiled form (C <= 256), the one the launcher prefers.  Per (tile of 32 key pixels, chunk of planes) the gradient with
// respect to the correlation matrix restricted to the chunk's band,
//   dG[p][t] = sum_{s,k : t_k(p,s) = t} g'(p,s) w_k(p,s),
// is built in shared memory over the bounding box of the band (<= kBwdRMax texels), and both gradients become small dense
// contractions over it:  d key[p][c] += sum_t dG[p][t] src[t][c]   and   d src[t][c] += sum_p dG[p][t] key[p][c].
// A source texel is read once and reduced into d src once per (tile, chunk) instead of once per (pixel, run, corner):
// ~12x fewer bytes through the L2 reduction path at the KITTI shape.  Thread = channel: 32 key values and 32 key-gradient
// accumulators in registers, dG column (32 floats) broadcast from shared memory, texel rows of the channels-last maps
// coalesced.  A chunk whose box does not fit is retried with half the planes, a single plane with fewer pixels
// (one pixel x one plane = 4 texels always fits).
// ---------------------------------------------------------------------------------------------------------
constexpr int kBwdRMax = 512;   // texels of a band box (72 KB of dG, texel list and flags)
constexpr int kBwdPlanes = 128; // planes per chunk to start from (measured: 32 -> 0.89, 64 -> 0.69, 128 -> 0.63 ms at cfg2)

struct BwdSample {
  bool valid;
  int x0c, y0c, dx, dy;
  float nw, ne, sw, se;  // weights times g'
};

__device__ __forceinline__ BwdSample bwd_sample(const PsweepBwdArgs& A, const Epi& E, int n, int v, int p, int s, int hs, int ws,
                                                float scale, size_t g_base, bool need_g) {
  BwdSample r;
  r.valid = false;
  const int P = A.h * A.w;
  if (p >= P) return r;
  const int yi = p / A.w, xi = p - yi * A.w;
  const EpiPix ep = epi_pixel(E, xi, yi);
  const float d = A.invd[n * A.ist[0] + s * A.ist[1] + yi * A.ist[2] + xi * A.ist[3]];
  float us, vs;
  bool geom;
  epi_sample(E, ep, d, us, vs, geom);
  const Bilinear b = bilinear_setup(pix_to_unnorm(us, ws), pix_to_unnorm(vs, hs), hs, ws);
  r.valid = sampling_mask(b.msum) && geom;
  if (!r.valid) return r;
  const int x0 = sat_index(b.x0f, ws), y0 = sat_index(b.y0f, hs);
  r.x0c = clampi(x0, 0, ws - 1);
  r.y0c = clampi(y0, 0, hs - 1);
  r.dx = clampi(x0 + 1, 0, ws - 1) - r.x0c;
  r.dy = clampi(y0 + 1, 0, hs - 1) - r.y0c;
  if (need_g) {
    const float g = fdiv(A.gcorr[(g_base + A.out_s0 + (s - A.s_begin)) * P + p], scale);
    r.nw = b.nw * g; r.ne = b.ne * g; r.sw = b.sw * g; r.se = b.se * g;
  }
  return r;
}

__global__ void __launch_bounds__(256, 2) psweep_corr_bwd_tile_kernel(const PsweepBwdArgs A) {
  extern __shared__ __align__(16) float smem[];
  float* s_dg = smem;                                        // [rmax][32]  dG, texel-major
  int* s_flag = reinterpret_cast<int*>(s_dg + A.rmax * 32);    // [rmax] texel touched
  int2* s_list = reinterpret_cast<int2*>(s_flag + A.rmax);     // [rmax] touched texels: (box index, offset in the source map)
  __shared__ int s_box[4];                                   // xmin, xmax, ymin, ymax
  __shared__ int s_count;

  const int v = blockIdx.z, n = blockIdx.y;
  const int C = A.C, P = A.h * A.w;
  const int p0 = (blockIdx.x / A.zsplit) * kTilePx;
  const int zlen = (A.s_end - A.s_begin + A.zsplit - 1) / A.zsplit;
  const int z_begin = A.s_begin + (blockIdx.x % A.zsplit) * zlen;
  const int z_end = min(A.s_end, z_begin + zlen);
  const int t = threadIdx.x, lane = t & 31;
  const int hs = A.hs[v], ws = A.ws[v];
  const float* __restrict__ src = A.src[v] + (size_t)n * hs * ws * C;
  float* __restrict__ gsrc = A.gsrc[v] ? A.gsrc[v] + (size_t)n * hs * ws * C : nullptr;
  const bool ch = t < C;        // this thread owns channel t
  const bool do_key = A.gkey != nullptr;

  float key[32], gk[32];
  {
    const float* kp = A.key + ((size_t)n * C + (ch ? t : 0)) * P + p0;
#pragma unroll
    for (int p = 0; p < 32; ++p) {
      key[p] = (ch && p0 + p < P) ? __ldg(kp + p) : 0.f;
      gk[p] = 0.f;
    }
  }
  const Epi E = epi_from_calib(A.Kk + n * 9, A.Ks[v] + n * 9, A.T[v] + n * 16, A.h, A.w, hs, ws);
  const float scale = (A.norm == RMVD_NORM_DIM) ? sqrtf((float)C) : 1.0f;
  const size_t g_base = ((size_t)(A.v0 + v) * A.N + n) * A.out_S;

  // bounding box of the clamped corners of the valid samples of (planes [sc, sc+ns), pixels [pa, pa+npx)); false: no sample
  auto find_box = [&](int sc, int ns, int pa, int npx) -> bool {
    __syncthreads();
    if (t == 0) { s_box[0] = 1 << 30; s_box[1] = -1; s_box[2] = 1 << 30; s_box[3] = -1; }
    __syncthreads();
    int xlo = 1 << 30, xhi = -1, ylo = 1 << 30, yhi = -1;
    for (int idx = t; idx < npx * ns; idx += 256) {
      const int q = idx % npx, sp = idx / npx;
      const BwdSample r = bwd_sample(A, E, n, v, p0 + pa + q, sc + sp, hs, ws, scale, g_base, false);
      if (r.valid) {
        xlo = min(xlo, r.x0c); xhi = max(xhi, r.x0c + r.dx);
        ylo = min(ylo, r.y0c); yhi = max(yhi, r.y0c + r.dy);
      }
    }
    xlo = __reduce_min_sync(kFullMask, xlo); xhi = __reduce_max_sync(kFullMask, xhi);
    ylo = __reduce_min_sync(kFullMask, ylo); yhi = __reduce_max_sync(kFullMask, yhi);
    if (lane == 0 && xhi >= 0) {
      atomicMin(&s_box[0], xlo); atomicMax(&s_box[1], xhi);
      atomicMin(&s_box[2], ylo); atomicMax(&s_box[3], yhi);
    }
    __syncthreads();
    return s_box[1] >= 0;
  };

  // dG over the box, then the two contractions
  auto process = [&](int sc, int ns, int pa, int npx) {
    const int bx = s_box[0], by = s_box[2];
    const int bw = s_box[1] - bx + 1, bh = s_box[3] - by + 1;
    const int R = bw * bh;
    for (int i = t; i < R * 32; i += 256) s_dg[i] = 0.f;
    for (int i = t; i < R; i += 256) s_flag[i] = 0;
    if (t == 0) s_count = 0;
    __syncthreads();
    for (int idx = t; idx < npx * ns; idx += 256) {
      const int q = idx % npx, sp = idx / npx;
      const BwdSample r = bwd_sample(A, E, n, v, p0 + pa + q, sc + sp, hs, ws, scale, g_base, true);
      if (r.valid) {
        const int col = pa + q;
        const int t_nw = (r.y0c - by) * bw + (r.x0c - bx);
        const int t_ne = t_nw + r.dx, t_sw = t_nw + r.dy * bw, t_se = t_sw + r.dx;
        if (r.nw != 0.f) { atomicAdd(&s_dg[t_nw * 32 + col], r.nw); s_flag[t_nw] = 1; }
        if (r.ne != 0.f) { atomicAdd(&s_dg[t_ne * 32 + col], r.ne); s_flag[t_ne] = 1; }
        if (r.sw != 0.f) { atomicAdd(&s_dg[t_sw * 32 + col], r.sw); s_flag[t_sw] = 1; }
        if (r.se != 0.f) { atomicAdd(&s_dg[t_se * 32 + col], r.se); s_flag[t_se] = 1; }
      }
    }
    __syncthreads();
    for (int i0 = 0; i0 < R; i0 += 256) {  // compact the touched texels of the box into a list (one shared atomic per warp)
      const int i = i0 + t;
      const bool on = i < R && s_flag[i] != 0;
      const unsigned bal = __ballot_sync(kFullMask, on);
      int base = 0;
      if (lane == 0 && bal) base = atomicAdd(&s_count, __popc(bal));
      base = __shfl_sync(kFullMask, base, 0);
      if (on) {
        const int ty = i / bw, tx = i - ty * bw;
        s_list[base + __popc(bal & ((1u << lane) - 1u))] = make_int2(i, (by + ty) * ws + bx + tx);
      }
    }
    __syncthreads();
    if (ch) {
      const int cnt = s_count;  // touched texels of the box
      int2 e = cnt > 0 ? s_list[0] : make_int2(0, 0);
      float sv = (do_key && cnt > 0) ? __ldg(src + (size_t)e.y * C + t) : 0.f;
      for (int i = 0; i < cnt; ++i) {
        const int2 en = (i + 1 < cnt) ? s_list[i + 1] : e;
        const float svn = do_key ? __ldg(src + (size_t)en.y * C + t) : 0.f;  // next texel's value under this one's FMAs
        const float4* dg = reinterpret_cast<const float4*>(s_dg + e.x * 32);
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);   // four independent chains
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const float4 d = dg[k];
          gk[4 * k + 0] = fmaf(d.x, sv, gk[4 * k + 0]); acc.x = fmaf(d.x, key[4 * k + 0], acc.x);
          gk[4 * k + 1] = fmaf(d.y, sv, gk[4 * k + 1]); acc.y = fmaf(d.y, key[4 * k + 1], acc.y);
          gk[4 * k + 2] = fmaf(d.z, sv, gk[4 * k + 2]); acc.z = fmaf(d.z, key[4 * k + 2], acc.z);
          gk[4 * k + 3] = fmaf(d.w, sv, gk[4 * k + 3]); acc.w = fmaf(d.w, key[4 * k + 3], acc.w);
        }
        if (gsrc) atomicAdd(gsrc + (size_t)e.y * C + t, (acc.x + acc.y) + (acc.z + acc.w));  // RED, 128-byte rows per warp
        e = en;
        sv = svn;
      }
    }
  };

  int ns_try = A.planes0;
  const long long rmax = A.rmax;
  int sc = z_begin;
  while (sc < z_end) {
    const int ns = min(ns_try, z_end - sc);
    if (!find_box(sc, ns, 0, kTilePx)) { sc += ns; continue; }                      // no valid sample in the chunk
    const long long area = (long long)(s_box[1] - s_box[0] + 1) * (s_box[3] - s_box[2] + 1);
    if (area <= rmax) {
      process(sc, ns, 0, kTilePx);
      sc += ns;
      continue;
    }
    if (ns > 1) { ns_try = ns / 2; continue; }                                      // retry with half the planes
    for (int pa = 0; pa < kTilePx;) {                                               // one plane: split the pixels
      int npx = min(16, kTilePx - pa);
      bool any;
      for (;;) {
        any = find_box(sc, 1, pa, npx);
        if (!any || (long long)(s_box[1] - s_box[0] + 1) * (s_box[3] - s_box[2] + 1) <= rmax || npx == 1) break;
        npx >>= 1;
      }
      if (any) process(sc, 1, pa, npx);
      pa += npx;
    }
    sc += 1;
  }

  if (do_key && ch) {  // several views (CTAs of grid.z) add into the zeroed NCHW key gradient
    float* gp = A.gkey + ((size_t)n * C + t) * P + p0;
#pragma unroll
    for (int p = 0; p < 32; ++p)
      if (p0 + p < P && gk[p] != 0.f) atomicAdd(gp + p, gk[p]);
  }
}

// channels-last [N,P,C] <-> NCHW [N,C,P] transposes (32 x 32 tiles through shared memory)
__global__ void __launch_bounds__(256) bwd_nchw_to_nhwc_kernel(const float* __restrict__ src, float* __restrict__ dst,
                                                               int C, int P) {
  __shared__ float tile[32][33];
  const int n = blockIdx.z, p0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const float* s = src + (size_t)n * C * P;
  float* d = dst + (size_t)n * P * C;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int c = c0 + ty + 8 * k;
    tile[ty + 8 * k][tx] = (c < C && p0 + tx < P) ? s[(size_t)c * P + p0 + tx] : 0.f;
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int p = p0 + ty + 8 * k, c = c0 + tx;
    if (p < P && c < C) d[(size_t)p * C + c] = tile[tx][ty + 8 * k];
  }
}
__global__ void __launch_bounds__(256) bwd_nhwc_to_nchw_kernel(const float* __restrict__ src, float* __restrict__ dst,
                                                               int C, int P) {
  __shared__ float tile[32][33];
  const int n = blockIdx.z, p0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const float* s = src + (size_t)n * P * C;
  float* d = dst + (size_t)n * C * P;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int p = p0 + ty + 8 * k;
    tile[ty + 8 * k][tx] = (p < P && c0 + tx < C) ? s[(size_t)p * C + c0 + tx] : 0.f;
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int c = c0 + ty + 8 * k, p = p0 + tx;
    if (p < P && c < C) d[(size_t)c * P + p] = tile[tx][ty + 8 * k];
  }
}

static size_t slot_bytes(int N, int C, long long P) { return (((size_t)N * C * P * sizeof(float)) + 255) / 256 * 256; }

template <int kC4>
static int launch_bwd(const PsweepBwdArgs& A, int nviews, cudaStream_t st) {
  const size_t smem = (size_t)kTilePx * (A.C + 4) * sizeof(float) * (kC4 == 0 ? 2 : 1);
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(psweep_corr_bwd_kernel<kC4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
  }
  const int P = A.h * A.w;
  dim3 grid((P + kTilePx - 1) / kTilePx, A.N, nviews);
  psweep_corr_bwd_kernel<kC4><<<grid, kWarps * 32, smem, st>>>(A);
  return (int)cudaGetLastError();
}

}  // namespace rmvd

using namespace rmvd;

extern "C" size_t rmvd_psweep_bwd_workspace_bytes(int N, int C, int h, int w, int V, const int* hs, const int* ws) {
  size_t tot = slot_bytes(N, C, (long long)h * w);                                      // key gradient, channels-last
  for (int v = 0; v < V; ++v) tot += 2 * slot_bytes(N, C, (long long)hs[v] * ws[v]);    // source copy + its gradient
  return tot;
}

extern "C" int rmvd_psweep_corr_bwd(const float* feat_key, const float* const* feat_src, const float* intrinsics_key,
                                    const float* const* intrinsics_src, const float* const* src_to_key,
                                    const float* invdepths, const int64_t* inv_stride, int N, int C, int h, int w,
                                    int V, const int* hs, const int* ws, int S, int s_begin, int s_end,
                                    int normalize_mode, const float* grad_corr, int out_S, int out_s0,
                                    float* grad_key, float* const* grad_src, void* workspace, size_t workspace_bytes,
                                    void* stream) {
  if (!feat_key || !feat_src || !intrinsics_key || !intrinsics_src || !src_to_key || !invdepths || !inv_stride ||
      !hs || !ws || !grad_corr || !grad_src)
    return RMVD_E_NULL;
  if (N <= 0 || C <= 0 || h <= 0 || w <= 0 || V <= 0 || S <= 0 || out_S <= 0 || N > 65535) return RMVD_E_SHAPE;
  if (s_begin < 0 || s_end > S || s_begin > s_end || out_s0 < 0 || out_s0 + (s_end - s_begin) > out_S)
    return RMVD_E_RANGE;
  // L2-normalised inputs (normalize=True / "before") are differentiated by the caller around a NORM_NONE call
  if (normalize_mode != RMVD_NORM_NONE && normalize_mode != RMVD_NORM_DIM) return RMVD_E_MODE;
  for (int v = 0; v < V; ++v) {
    if (!feat_src[v] || !intrinsics_src[v] || !src_to_key[v]) return RMVD_E_NULL;
    if (hs[v] <= 0 || ws[v] <= 0 || hs[v] >= 32768 || ws[v] >= 32768) return RMVD_E_SHAPE;
  }
  if (!workspace || workspace_bytes < rmvd_psweep_bwd_workspace_bytes(N, C, h, w, V, hs, ws)) return RMVD_E_WORKSPACE;
  cudaStream_t st = (cudaStream_t)stream;
  const int P = h * w;
  const unsigned cb = (unsigned)((C + 31) / 32);

  // kernel choice: tiled form for C <= 256; RMVD_PSWEEP_BWD_KERNEL=direct|tile restricts it (A/B runs only)
  const char* force = getenv("RMVD_PSWEEP_BWD_KERNEL");
  const bool tile = (C <= 256) && !(force && strcmp(force, "direct") == 0);

  char* wsp = (char*)workspace;
  float* gkey_cl = (float*)wsp;
  wsp += slot_bytes(N, C, P);
  cudaError_t e;
  if (grad_key) {  // the tiled kernel reduces straight into the NCHW output, the direct one into a channels-last copy
    e = cudaMemsetAsync(tile ? grad_key : gkey_cl, 0, (size_t)N * C * P * sizeof(float), st);
    if (e != cudaSuccess) return (int)e;
  }
  for (int v0 = 0; v0 < V; v0 += kMaxViews) {
    const int nv = (V - v0 < kMaxViews) ? (V - v0) : kMaxViews;
    PsweepBwdArgs A{};
    A.key = feat_key; A.Kk = intrinsics_key; A.invd = invdepths; A.gcorr = grad_corr;
    A.gkey = grad_key ? (tile ? grad_key : gkey_cl) : nullptr;
    for (int k = 0; k < 4; ++k) A.ist[k] = inv_stride[k];
    float* g_cl[kMaxViews];
    for (int j = 0; j < nv; ++j) {
      const int v = v0 + j;
      const int Ps = hs[v] * ws[v];
      float* s_cl = (float*)wsp;
      wsp += slot_bytes(N, C, Ps);
      g_cl[j] = (float*)wsp;
      wsp += slot_bytes(N, C, Ps);
      dim3 g((Ps + 31) / 32, cb, N);
      bwd_nchw_to_nhwc_kernel<<<g, 256, 0, st>>>(feat_src[v], s_cl, C, Ps);
      if ((e = cudaGetLastError()) != cudaSuccess) return (int)e;
      if (grad_src[v]) {
        e = cudaMemsetAsync(g_cl[j], 0, (size_t)N * C * Ps * sizeof(float), st);
        if (e != cudaSuccess) return (int)e;
      }
      A.src[j] = s_cl;
      A.gsrc[j] = grad_src[v] ? g_cl[j] : nullptr;
      A.Ks[j] = intrinsics_src[v];
      A.T[j] = src_to_key[v];
      A.hs[j] = hs[v];
      A.ws[j] = ws[v];
    }
    A.N = N; A.C = C; A.h = h; A.w = w; A.S = S; A.s_begin = s_begin; A.s_end = s_end;
    A.norm = normalize_mode; A.out_S = out_S; A.out_s0 = out_s0; A.v0 = v0;
    int rc;
    if (tile) {
      // RMVD_BWD_PLANES / RMVD_BWD_RMAX: tuning overrides (A/B runs only)
      const char* ep = getenv("RMVD_BWD_PLANES");
      const char* er = getenv("RMVD_BWD_RMAX");
      A.planes0 = ep ? atoi(ep) : kBwdPlanes;
      A.rmax = (er ? atoi(er) : kBwdRMax) & ~1;   // even: the texel list behind the flags is 8-byte aligned
      if (A.planes0 < 1 || A.rmax < 4 || A.rmax > 1600) return RMVD_E_MODE;
      const size_t smem = (size_t)A.rmax * 35 * sizeof(float);
      e = cudaFuncSetAttribute(psweep_corr_bwd_tile_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (e != cudaSuccess) return (int)e;
      // CTAs per tile
      int dev = 0, sms = 148;
      cudaGetDevice(&dev);
      cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
      const long long tiles = (long long)((P + kTilePx - 1) / kTilePx) * N * nv, slots = 2LL * sms;
      const char* ez = getenv("RMVD_BWD_ZSPLIT");
      // fewer than ~3 waves of tiles leave the last wave ragged: halve / third the plane range per CTA (measured at cfg2, one
      // frame = 1.6 waves: 1 -> 0.68, 2 -> 0.62, 3 -> 0.66, 4 -> 0.68, 8 -> 0.87 ms; every split re-reads the tile-wide band)
      int zs = 1;
      while (zs < 3 && tiles * zs < 3 * slots && (s_end - s_begin) / (zs + 1) >= 64) ++zs;
      A.zsplit = ez ? atoi(ez) : zs;
      if (A.zsplit < 1 || A.zsplit > 64) return RMVD_E_MODE;
      dim3 grid((unsigned)(((P + kTilePx - 1) / kTilePx) * A.zsplit), N, nv);
      psweep_corr_bwd_tile_kernel<<<grid, 256, smem, st>>>(A);
      rc = (int)cudaGetLastError();
    } else
    switch (C) {
      case 128: rc = launch_bwd<1>(A, nv, st); break;
      case 256: rc = launch_bwd<2>(A, nv, st); break;
      case 384: rc = launch_bwd<3>(A, nv, st); break;
      case 512: rc = launch_bwd<4>(A, nv, st); break;
      default: rc = launch_bwd<0>(A, nv, st); break;
    }
    if (rc) return rc;
    for (int j = 0; j < nv; ++j) {
      const int v = v0 + j;
      if (!grad_src[v]) continue;
      const int Ps = hs[v] * ws[v];
      dim3 g((Ps + 31) / 32, cb, N);
      bwd_nhwc_to_nchw_kernel<<<g, 256, 0, st>>>(g_cl[j], grad_src[v], C, Ps);
      if ((e = cudaGetLastError()) != cudaSuccess) return (int)e;
    }
  }
  if (grad_key && !tile) {
    dim3 g((P + 31) / 32, cb, N);
    bwd_nhwc_to_nchw_kernel<<<g, 256, 0, st>>>(gkey_cl, grad_key, C, P);
    if ((e = cudaGetLastError()) != cudaSuccess) return (int)e;
  }
  return RMVD_OK;
}
