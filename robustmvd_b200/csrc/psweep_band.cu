// Banded-GEMM plane-sweep correlation kernel (sm_100a).
//
// Idea (the reference's own, restricted to where it matters): TorchCorr forms the all-pairs matrix
// G = key^T src (planesweep_corr.py:174) and bilinearly samples it (:181-186).  For one tile of 32 adjacent key
// pixels only the source texels under the tile's epipolar band are ever sampled (a few hundred of the hs*ws), and
// consecutive depth planes hit the same texels again and again.  So per CTA (32 key pixels x a chunk of planes):
//   phase 1  every (pixel, plane) sample is located with the bit-exact coordinate chain; the source texels its
//            four corners touch are marked in a shared-memory bitmap; a popcount scan compacts them to a list
//   phase 2  G[32 px][R texels] = key_tile . texel_list^T, an fp32 register-tiled GEMM (thread tile 8 px x 4 texels,
//            K = C in chunks of 32 staged with cp.async double buffering) -> G lives in shared memory only
//   phase 3  every sample is located again, its corners are looked up in G by bitmap rank and interpolated;
//            corr and mask rows (32 px = 128 B) are stored coalesced.
// FLOPs drop from 4*C per sample to C per DISTINCT texel of the band, and every source texel is fetched once per
// CTA.  If a chunk's band exceeds the G capacity the chunk is split (256 -> 64 -> 16 planes), and a 16-plane chunk
// that still does not fit (no reuse to exploit) is evaluated directly.
//
// Both feature maps are read from channels-last copies [N, pixels, C] (see nchw_to_nhwc_kernel).
#include "psweep.cuh"

namespace rmvd {

constexpr int kBT = 128;           // threads per CTA
constexpr int kRMax = 448;         // capacity of G in texels
constexpr int kGPitch = kRMax + 2; // kGPitch + 1 is odd: adjacent pixels reading adjacent texels hit distinct banks
constexpr int kKC = 32;            // channels per staged chunk
constexpr int kRowPitch = kKC + 4; // 36 floats: conflict-free 128-bit shared loads
constexpr int kTN = 128;           // texels per GEMM pass
constexpr int kMaxSrcTexels = 65536;

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gmem_src) {
  const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(d), "l"(gmem_src));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int kN>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(kN)); }

// One located sample: clamped corner texels, weights (0 for out-of-range corners), validity.
struct Located {
  int t_nw, t_ne, t_sw, t_se;  // texel indices y*ws + x of the clamped corners
  float nw, ne, sw, se;
  int x0, y0;                  // saturated floor indices (for the debug outputs)
  bool valid, geom;
};

__device__ __forceinline__ Located locate(const PsweepArgs& A, const Epi& E, int n, int s, int p, int hs, int ws) {
  Located L;
  const int yi = p / A.w, xi = p - yi * A.w;
  const EpiPix ep = epi_pixel(E, xi, yi);
  const float d = A.invd[n * A.ist[0] + s * A.ist[1] + yi * A.ist[2] + xi * A.ist[3]];
  float us, vs;
  epi_sample(E, ep, d, us, vs, L.geom);
  const Bilinear b = bilinear_setup(pix_to_unnorm(us, ws), pix_to_unnorm(vs, hs), hs, ws);
  L.valid = sampling_mask(b.msum) && L.geom;
  L.nw = b.nw; L.ne = b.ne; L.sw = b.sw; L.se = b.se;
  L.x0 = sat_index(b.x0f, ws);
  L.y0 = sat_index(b.y0f, hs);
  const int x0c = clampi(L.x0, 0, ws - 1), x1c = clampi(L.x0 + 1, 0, ws - 1);
  const int y0c = clampi(L.y0, 0, hs - 1), y1c = clampi(L.y0 + 1, 0, hs - 1);
  L.t_nw = y0c * ws + x0c; L.t_ne = y0c * ws + x1c; L.t_sw = y1c * ws + x0c; L.t_se = y1c * ws + x1c;
  return L;
}

struct BandSmem {
  float* G;        // [32][kGPitch]
  float* A;        // [2][32][kRowPitch]
  float* B;        // [2][kTN][kRowPitch]
  int* list;       // [kRMax + kTN] texel index of each G column
  unsigned* bits;  // [nwords] needed-texel bitmap over the source map
  int* prefix;     // [nwords] exclusive popcount scan of bits
  int* misc;       // [8] scratch: [0] = R, [1..4] warp sums
};

__device__ __forceinline__ int rank_of(const BandSmem& S, int tex) {
  const int wd = tex >> 5;
  return S.prefix[wd] + __popc(S.bits[wd] & ((1u << (tex & 31)) - 1u));
}

// phase 1: mark + compact.  Returns R (number of distinct texels) to every thread.
__device__ int band_mark(const PsweepArgs& A, const Epi& E, const BandSmem& S, int n, int p0, int P, int s0, int len,
                         int hs, int ws, int nwords) {
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  for (int i = t; i < nwords; i += kBT) S.bits[i] = 0u;
  __syncthreads();
  const int p = p0 + lane;
  if (p < P) {
    for (int pl = warp; pl < len; pl += kBT / 32) {
      const Located L = locate(A, E, n, s0 + pl, p, hs, ws);
      if (L.valid) {
        atomicOr(&S.bits[L.t_nw >> 5], 1u << (L.t_nw & 31));
        atomicOr(&S.bits[L.t_ne >> 5], 1u << (L.t_ne & 31));
        atomicOr(&S.bits[L.t_sw >> 5], 1u << (L.t_sw & 31));
        atomicOr(&S.bits[L.t_se >> 5], 1u << (L.t_se & 31));
      }
    }
  }
  __syncthreads();
  // exclusive scan of popcounts: each thread owns a contiguous run of words
  const int wpt = (nwords + kBT - 1) / kBT;
  const int w0 = t * wpt, w1 = min(nwords, w0 + wpt);
  int local = 0;
  for (int i = w0; i < w1; ++i) local += __popc(S.bits[i]);
  int incl = local;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int v = __shfl_up_sync(kFullMask, incl, o);
    if (lane >= o) incl += v;
  }
  if (lane == 31) S.misc[1 + warp] = incl;
  __syncthreads();
  int base = incl - local;
  for (int k = 0; k < warp; ++k) base += S.misc[1 + k];
  const int R = S.misc[1] + S.misc[2] + S.misc[3] + S.misc[4];
  for (int i = w0; i < w1; ++i) {
    S.prefix[i] = base;
    unsigned b = S.bits[i];
    if (R <= kRMax) {
      while (b) {
        const int j = __ffs(b) - 1;
        b &= b - 1;
        S.list[base++] = i * 32 + j;
      }
    } else {
      base += __popc(b);
    }
  }
  __syncthreads();
  return R;
}

// phase 2: G[32][R] = key_tile[32][C] . src[list[r]][C]^T
__device__ void band_gemm(const BandSmem& S, const float* __restrict__ key_nhwc, const float* __restrict__ src, int C,
                          int p0, int P, int R) {
  const int t = threadIdx.x, lane = t & 31, pg = t >> 5;
  const int nkc = C / kKC;
  const int npass = (R + kTN - 1) / kTN;
  const int T = npass * nkc;

  auto stage_load = [&](int it) {
    const int pass = it / nkc, kc = it - pass * nkc;
    float* sa = S.A + (it & 1) * 32 * kRowPitch;
    float* sb = S.B + (it & 1) * kTN * kRowPitch;
#pragma unroll
    for (int q = 0; q < 2; ++q) {  // A: 32 px x 8 pieces of 16 B
      const int id = t + kBT * q, px = id >> 3, piece = id & 7;
      const int p = min(p0 + px, P - 1);
      cp_async16(sa + px * kRowPitch + piece * 4, key_nhwc + (size_t)p * C + kc * kKC + piece * 4);
    }
#pragma unroll
    for (int q = 0; q < 8; ++q) {  // B: 128 texels x 8 pieces; 8 consecutive threads fetch one texel's 128 B
      const int id = t + kBT * q, row = id >> 3, piece = id & 7;
      const int r = pass * kTN + row;
      const int tex = S.list[r < R ? r : 0];
      cp_async16(sb + row * kRowPitch + piece * 4, src + (size_t)tex * C + kc * kKC + piece * 4);
    }
    cp_async_commit();
  };

  float acc[8][4];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  stage_load(0);
  for (int it = 0; it < T; ++it) {
    if (it + 1 < T) {
      stage_load(it + 1);
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    __syncthreads();
    const float* sa = S.A + (it & 1) * 32 * kRowPitch + pg * 8 * kRowPitch;
    const float* sb = S.B + (it & 1) * kTN * kRowPitch + lane * kRowPitch;
#pragma unroll
    for (int kq = 0; kq < kKC / 4; ++kq) {
      float4 a[8], b[4];
#pragma unroll
      for (int i = 0; i < 8; ++i) a[i] = *reinterpret_cast<const float4*>(sa + i * kRowPitch + kq * 4);
#pragma unroll
      for (int j = 0; j < 4; ++j) b[j] = *reinterpret_cast<const float4*>(sb + j * 32 * kRowPitch + kq * 4);
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = dot4(a[i], b[j], acc[i][j]);
    }
    const int pass = it / nkc, kc = it - pass * nkc;
    if (kc == nkc - 1) {
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int r = pass * kTN + j * 32 + lane;
          if (r < kRMax) S.G[(pg * 8 + i) * kGPitch + r] = acc[i][j];
          acc[i][j] = 0.f;
        }
    }
    __syncthreads();
  }
}

// phase 3: interpolate G and store.  Warp = one plane x 32 pixels (one 128-byte row of corr and of mask).
__device__ void band_sample(const PsweepArgs& A, const Epi& E, const BandSmem& S, int n, int v, int p0, int P, int s0,
                            int len, int hs, int ws, float scale) {
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int p = p0 + lane;
  if (p >= P) return;
  const size_t out_base = ((size_t)(A.v0 + v) * A.N + n) * A.out_S;
  const float* g = S.G + lane * kGPitch;
  for (int pl = warp; pl < len; pl += kBT / 32) {
    const int s = s0 + pl;
    const Located L = locate(A, E, n, s, p, hs, ws);
    float val = 0.f;
    if (L.valid) {
      const float g_nw = g[rank_of(S, L.t_nw)], g_ne = g[rank_of(S, L.t_ne)];
      const float g_sw = g[rank_of(S, L.t_sw)], g_se = g[rank_of(S, L.t_se)];
      val = fdiv(g_nw * L.nw + g_ne * L.ne + g_sw * L.sw + g_se * L.se, scale);
    }
    const size_t o = (out_base + A.out_s0 + (s - A.s_begin)) * P + p;
    A.corr[o] = val;
    A.mask[o] = L.valid ? 1.f : 0.f;
    if (A.dbg_ix0) {
      A.dbg_ix0[o] = L.x0;
      A.dbg_iy0[o] = L.y0;
      if (A.dbg_geom) A.dbg_geom[o] = L.geom ? 1 : 0;
    }
  }
}

// direct evaluation of a small chunk whose band does not fit G (no texel reuse to exploit)
__device__ void band_direct(const PsweepArgs& A, const Epi& E, const float* __restrict__ key_nhwc,
                            const float* __restrict__ src, int n, int v, int p0, int P, int s0, int len, int hs, int ws,
                            float scale) {
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int C = A.C;
  const int p = p0 + lane;
  const size_t out_base = ((size_t)(A.v0 + v) * A.N + n) * A.out_S;
  for (int pl = warp; pl < len; pl += kBT / 32) {
    const int s = s0 + pl;
    Located L{};
    if (p < P) L = locate(A, E, n, s, p, hs, ws);
    unsigned vm = __ballot_sync(kFullMask, p < P && L.valid);
    float res = 0.f;
    while (vm) {
      const int j = __ffs(vm) - 1;
      vm &= vm - 1;
      const int t0 = __shfl_sync(kFullMask, L.t_nw, j), t1 = __shfl_sync(kFullMask, L.t_ne, j);
      const int t2 = __shfl_sync(kFullMask, L.t_sw, j), t3 = __shfl_sync(kFullMask, L.t_se, j);
      const float nw = __shfl_sync(kFullMask, L.nw, j), ne = __shfl_sync(kFullMask, L.ne, j);
      const float sw = __shfl_sync(kFullMask, L.sw, j), se = __shfl_sync(kFullMask, L.se, j);
      const float* kr = key_nhwc + (size_t)(p0 + j) * C;
      float a = 0.f;
      for (int c = lane * 4; c < C; c += 128) {
        const float4 r = ldg4(kr + c);
        const float4 q0 = ldg4(src + (size_t)t0 * C + c), q1 = ldg4(src + (size_t)t1 * C + c);
        const float4 q2 = ldg4(src + (size_t)t2 * C + c), q3 = ldg4(src + (size_t)t3 * C + c);
        a += dot4(r, q0, 0.f) * nw + dot4(r, q1, 0.f) * ne + dot4(r, q2, 0.f) * sw + dot4(r, q3, 0.f) * se;
      }
      a = warp_sum(a);
      if (lane == j) res = a;
    }
    if (p < P) {
      const size_t o = (out_base + A.out_s0 + (s - A.s_begin)) * P + p;
      A.corr[o] = L.valid ? fdiv(res, scale) : 0.f;
      A.mask[o] = L.valid ? 1.f : 0.f;
      if (A.dbg_ix0) {
        A.dbg_ix0[o] = L.x0;
        A.dbg_iy0[o] = L.y0;
        if (A.dbg_geom) A.dbg_geom[o] = L.geom ? 1 : 0;
      }
    }
  }
}

__global__ void __launch_bounds__(kBT, 2) psweep_band_kernel(const PsweepArgs A, const float* __restrict__ key_nhwc_all) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int v = blockIdx.z, n = blockIdx.y;
  const int P = A.h * A.w, C = A.C;
  const int p0 = blockIdx.x * kTilePx;
  const int hs = A.hs[v], ws = A.ws[v];
  const int nwords = (hs * ws + 31) >> 5;
  BandSmem S;
  S.G = reinterpret_cast<float*>(smem_raw);
  S.A = S.G + 32 * kGPitch;
  S.B = S.A + 2 * 32 * kRowPitch;
  S.list = reinterpret_cast<int*>(S.B + 2 * kTN * kRowPitch);
  S.bits = reinterpret_cast<unsigned*>(S.list + kRMax + kTN);
  S.prefix = reinterpret_cast<int*>(S.bits + nwords);
  S.misc = S.prefix + nwords;

  const float* __restrict__ src = A.src[v] + (size_t)n * hs * ws * C;
  const float* __restrict__ key_nhwc = key_nhwc_all + (size_t)n * P * C;
  const Epi E = epi_from_calib(A.Kk + n * 9, A.Ks[v] + n * 9, A.T[v] + n * 16, A.h, A.w, hs, ws);
  const float scale = (A.norm == RMVD_NORM_DIM) ? sqrtf((float)C) : 1.0f;

  int s = A.s_begin;
  int len = min(256, A.s_end - s);
  while (s < A.s_end) {
    len = min(len, A.s_end - s);
    const int R = band_mark(A, E, S, n, p0, P, s, len, hs, ws, nwords);
    if (R <= kRMax) {
      if (R > 0) band_gemm(S, key_nhwc, src, C, p0, P, R);
      band_sample(A, E, S, n, v, p0, P, s, len, hs, ws, scale);
      s += len;
    } else if (len > 16) {
      len = (len + 3) / 4;
    } else {
      band_direct(A, E, key_nhwc, src, n, v, p0, P, s, len, hs, ws, scale);
      s += len;
    }
    __syncthreads();
  }
}

static size_t band_smem_bytes(int max_src_texels) {
  const size_t nwords = (max_src_texels + 31) / 32;
  return (size_t)(32 * kGPitch + 2 * 32 * kRowPitch + 2 * kTN * kRowPitch) * 4 + (size_t)(kRMax + kTN) * 4 +
         nwords * 8 + 8 * 4;
}

int launch_psweep_band(const PsweepArgs& A, const float* key_nhwc, int nviews, cudaStream_t st) {
  if (A.C % kKC != 0) return -1000;
  int maxP = 0;
  for (int v = 0; v < nviews; ++v) maxP = max(maxP, A.hs[v] * A.ws[v]);
  if (maxP > kMaxSrcTexels) return -1000;
  const size_t smem = band_smem_bytes(maxP);
  cudaError_t e = cudaFuncSetAttribute(psweep_band_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return (int)e;
  const int P = A.h * A.w;
  dim3 grid((P + kTilePx - 1) / kTilePx, A.N, nviews);
  psweep_band_kernel<<<grid, kBT, smem, st>>>(A, key_nhwc);
  return (int)cudaGetLastError();
}

}  // namespace rmvd
