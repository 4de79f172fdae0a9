// MVSNet cost-volume path for sm_100a: homography warp fused with the multi-view variance, and the
// softmax / expected-depth / confidence regression in one pass.
//
// Replaces homo_warp + the volume assembly (blocks/utils.py:222-268, mvsnet.py:123-136) and
// softmax + depth_regression + 4-bin confidence (mvsnet.py:140-160, blocks/utils.py:271-274).
//
// Data layout in HBM: features [B,Vt,C,h,w] and cost volumes [B,D,h,w] exactly as the reference holds them;
// the variance volume [B,C,D,h,w] is written once with 128-byte coalesced rows along w; the warped source
// volumes (1.67 GB per view at DTU size) are never materialised.
#include <cstdlib>

#include "mvs_coord.cuh"

namespace rmvd {

struct Gather {
  int o_nw, o_ne, o_sw, o_se;
  float nw, ne, sw, se;
};
__device__ __forceinline__ Gather make_gather(float ix, float iy, int H, int W, int* dbg_x0 = nullptr, int* dbg_y0 = nullptr) {
  const Bilinear b = bilinear_setup(ix, iy, H, W);
  const int x0 = isnan(b.x0f) ? -2 : sat_index(b.x0f, W), y0 = isnan(b.y0f) ? -2 : sat_index(b.y0f, H);
  if (dbg_x0) { *dbg_x0 = x0; *dbg_y0 = y0; }
  const int x0c = clampi(x0, 0, W - 1), x1c = clampi(x0 + 1, 0, W - 1);
  const int y0c = clampi(y0, 0, H - 1), y1c = clampi(y0 + 1, 0, H - 1);
  Gather g;
  g.o_nw = y0c * W + x0c; g.o_ne = y0c * W + x1c; g.o_sw = y1c * W + x0c; g.o_se = y1c * W + x1c;
  g.nw = b.nw; g.ne = b.ne; g.sw = b.sw; g.se = b.se;
  return g;
}
__device__ __forceinline__ float gather_val(const float* __restrict__ img, const Gather& g) {
  float acc = fmul(__ldg(img + g.o_nw), g.nw);
  acc = fadd(acc, fmul(__ldg(img + g.o_ne), g.ne));
  acc = fadd(acc, fmul(__ldg(img + g.o_sw), g.sw));
  acc = fadd(acc, fmul(__ldg(img + g.o_se), g.se));
  return acc;
}

// ---------------------------------------------------------------------------------------------------------
// Fused variance volume.  Thread = one voxel (d, y, x) with x fastest; blockIdx.y = group of kCh channels.
// Per source view the thread derives its sampling footprint once and gathers kCh channels, keeping the running
// sum and sum of squares in registers (mvsnet.py:124-133); variance = sq/V - (sum/V)^2 (:135).
// ---------------------------------------------------------------------------------------------------------
//   kCvp: CVP-MVSNet's proj_cost (cvp_mvsnet_components.py:375-457): per-pixel depth hypotheses [B,D,h,w], the chain of
//   cvp_coord(), and the reference's in-place quirk -- `volume_sq_sum = volume_sum.pow_(2)` (:392) squares volume_sum
//   itself, so BOTH running sums start from ref^2: cost = (ref^2 + sum w^2)/V - ((ref^2 + sum w)/V)^2.
template <int kCh, bool kCvp = false>
__global__ void __launch_bounds__(256) mvs_variance_kernel(const float* __restrict__ feats, const float* __restrict__ proj,
                                                           const float* __restrict__ depths, int B, int Vt, int C,
                                                           int h, int w, int D, int d_begin, int d_end,
                                                           float* __restrict__ variance, int out_D, int out_d0) {
  __shared__ float s_M[kMaxMvsViews][12];
  const int b = blockIdx.z;
  const int P = h * w;
  const int c0 = blockIdx.y * kCh;
  // per-view transforms (view v >= 1): rows 0..2 of proj[b,v] @ proj[b,0]
  for (int e = threadIdx.x; e < (Vt - 1) * 12; e += blockDim.x) {
    const int v = 1 + e / 12, i = (e % 12) / 4, j = e % 4;
    s_M[v - 1][i * 4 + j] = matmul4_elem(proj + ((size_t)b * Vt + v) * 16, proj + ((size_t)b * Vt) * 16, i, j);
  }
  __syncthreads();
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const int Dl = d_end - d_begin;
  if (idx >= (long long)Dl * P) return;
  const int dl = (int)(idx / P), p = (int)(idx - (long long)dl * P);
  const int d = d_begin + dl;
  const int yi = p / w, xi = p - yi * w;
  const float depth = kCvp ? depths[((size_t)b * D + d) * P + p] : depths[(size_t)b * D + d];

  float sum[kCh], sq[kCh];
  const float* ref = feats + ((size_t)b * Vt) * C * P + p;
#pragma unroll
  for (int k = 0; k < kCh; ++k) {
    const float r = (c0 + k < C) ? __ldg(ref + (size_t)(c0 + k) * P) : 0.f;
    sum[k] = kCvp ? fmul(r, r) : r;   // volume_sum = ref_volume (:125); CVP: squared in place with volume_sq_sum
    sq[k] = fmul(r, r);               // ref_volume ** 2 (:126)
  }
  for (int v = 1; v < Vt; ++v) {
    float ix, iy;
    if (kCvp) cvp_coord(s_M[v - 1], xi, yi, depth, h, w, ix, iy);
    else homo_coord(s_M[v - 1], xi, yi, depth, h, w, ix, iy);
    const Gather g = make_gather(ix, iy, h, w);
    const float* src = feats + ((size_t)b * Vt + v) * C * P;
#pragma unroll
    for (int k = 0; k < kCh; ++k) {
      if (c0 + k < C) {
        const float val = gather_val(src + (size_t)(c0 + k) * P, g);
        sum[k] = fadd(sum[k], val);
        sq[k] = fadd(sq[k], fmul(val, val));
      }
    }
  }
  const float Vf = (float)Vt;
  float* out = variance + (((size_t)b * C + c0) * out_D + out_d0 + dl) * P + p;
#pragma unroll
  for (int k = 0; k < kCh; ++k) {
    if (c0 + k < C) {
      const float mean = fdiv(sum[k], Vf);
      out[(size_t)k * out_D * P] = fsub(fdiv(sq[k], Vf), fmul(mean, mean));  // :135
    }
  }
}

// ---------------------------------------------------------------------------------------------------------
// Fused variance volume, staged form.  CTA = 64 consecutive key pixels of one row x 4 consecutive depth planes,
// thread = one voxel x kCh channels (sum and sum of squares stay in registers across the views).  Per source view:
//   1. every thread locates its sample once with the exact chain (homo_coord + bilinear_setup);
//   2. the CTA reduces the bounding rectangle of all corners that carry weight -- exact, so nothing can fall outside;
//   3. the rectangle of all kCh channels is copied NCHW -> shared memory with coalesced row segments
//      ([channel][kVarCap] floats, row pitch = rectangle width), so each source element leaves L2 once per CTA
//      instead of once per voxel that touches it;
//   4. the 4 corners x kCh channels are read from shared memory: adjacent lanes read adjacent words.
// A view whose rectangle does not fit (strong rotation between the views) is gathered straight from global memory.
// Bound: 16 shared-memory words per output word -> at 128 B/clk/SM about a third of the HBM write roofline.
// ---------------------------------------------------------------------------------------------------------
constexpr int kVarPx = 64;       // key pixels per CTA (one row segment)
constexpr int kVarDB = 4;        // depth planes per CTA
constexpr int kVarCap = 768;     // staged floats per channel (rectangle area limit)
constexpr int kVarThreads = kVarPx * kVarDB;

template <int kCh, bool kFull>   // kFull: C is a multiple of kCh (no per-channel guards)
__global__ void __launch_bounds__(kVarThreads, 2) mvs_variance_tile_kernel(
    const float* __restrict__ feats, const float* __restrict__ proj, const float* __restrict__ depths, int B, int Vt, int C,
    int h, int w, int D, int d_begin, int d_end, float* __restrict__ variance, int out_D, int out_d0, int tiles_x, int ncg, int vec16) {
  extern __shared__ __align__(16) float s_tile[];   // [kCh][kVarCap]
  __shared__ float s_M[kMaxMvsViews][12];
  __shared__ int s_box[2][4];   // bounding rectangle {xlo, -xhi, ylo, -yhi} of the current view (min-reduced), two sets in turn
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int b = blockIdx.z;
  const int P = h * w;
  const int yi = blockIdx.x / tiles_x, bx = blockIdx.x - yi * tiles_x;
  const int dblk = blockIdx.y / ncg, cg = blockIdx.y - dblk * ncg;
  const int c0 = cg * kCh;
  for (int e = t; e < (Vt - 1) * 12; e += kVarThreads) {
    const int v = 1 + e / 12, i = (e % 12) / 4, j = e % 4;
    s_M[v - 1][i * 4 + j] = matmul4_elem(proj + ((size_t)b * Vt + v) * 16, proj + ((size_t)b * Vt) * 16, i, j);
  }
  if (t < 8) s_box[t >> 2][t & 3] = 0x7fffffff;
  __syncthreads();
  const float y_w = rcp_refined((float)(w - 1) * 0.5f), y_h = rcp_refined((float)(h - 1) * 0.5f);
  const int xi = bx * kVarPx + (t & (kVarPx - 1));
  const int dl = dblk * kVarDB + (t / kVarPx);
  const int d = d_begin + dl;
  const bool active = (xi < w) && (d < d_end);
  const int p = yi * w + min(xi, w - 1);
  const float depth = depths[(size_t)b * D + min(d, D - 1)];

  float sum[kCh], sq[kCh];
  {
    const float* ref = feats + (((size_t)b * Vt) * C + c0) * P + p;   // running pointer: one 64-bit add per channel
    const int nch = active ? (kFull ? kCh : min(kCh, C - c0)) : 0;
#pragma unroll
    for (int k = 0; k < kCh; ++k, ref += P) {
      const float r = (kFull ? active : (k < nch)) ? __ldg(ref) : 0.f;
      sum[k] = r;               // volume_sum = ref_volume (:125)
      sq[k] = r * r;            // ref_volume ** 2 (:126)
    }
  }
  for (int v = 1; v < Vt; ++v) {
    float ix, iy;
    homo_coord_shared(s_M[v - 1], xi, yi, depth, h, w, y_w, y_h, ix, iy);
    const Bilinear bl = bilinear_setup(ix, iy, h, w);
    const int x0 = isnan(bl.x0f) ? -2 : sat_index(bl.x0f, w), y0 = isnan(bl.y0f) ? -2 : sat_index(bl.y0f, h);
    const bool left = (bl.nw != 0.f) || (bl.sw != 0.f), right = (bl.ne != 0.f) || (bl.se != 0.f);
    const bool top = (bl.nw != 0.f) || (bl.ne != 0.f), bottom = (bl.sw != 0.f) || (bl.se != 0.f);
    const bool needed = active && (left || right);
    // bounding rectangle of the corners that carry weight (a non-zero weight implies the corner is inside the map)
    int xlo = needed ? (left ? x0 : x0 + 1) : 0x7fffffff, xhi = needed ? (right ? x0 + 1 : x0) : -0x7fffffff;
    int ylo = needed ? (top ? y0 : y0 + 1) : 0x7fffffff, yhi = needed ? (bottom ? y0 + 1 : y0) : -0x7fffffff;
    // warp reduction, then one shared atomic per warp and bound (lanes 0..3); maxima travel negated
    xlo = __reduce_min_sync(kFullMask, xlo); xhi = -__reduce_max_sync(kFullMask, xhi);
    ylo = __reduce_min_sync(kFullMask, ylo); yhi = -__reduce_max_sync(kFullMask, yhi);
    int* box = s_box[v & 1];
    if (lane < 4) atomicMin(box + lane, lane == 0 ? xlo : lane == 1 ? xhi : lane == 2 ? ylo : yhi);
    __syncthreads();
    xlo = box[0]; xhi = -box[1]; ylo = box[2]; yhi = -box[3];
    if (t < 4) s_box[(v + 1) & 1][t] = 0x7fffffff;          // next view's set: last read before this view's first barrier
    const float* src = feats + (((size_t)b * Vt + v) * C + c0) * P;
    if (xlo <= xhi) {                                       // uniform: someone samples inside the map
      // staged rectangle: columns [xo, xo + rw), rows [ylo, ylo + rh); with 16-byte copies the columns are widened to
      // multiples of 4 (w % 4 == 0 and a 16-byte aligned base keep every row segment aligned and inside its row)
      const int xo = vec16 ? (xlo & ~3) : xlo;
      const int rw = vec16 ? ((xhi - xo + 4) & ~3) : (xhi - xo + 1), rh = yhi - ylo + 1;
      if (rw * rh <= kVarCap) {                             // uniform
        const float* g0 = src + (size_t)ylo * w + xo;
        if (vec16) {
          const int cw = rw >> 2, nchunk = rh * cw;         // 16-byte chunks per row / per channel
          for (int q = t & 63; q < nchunk; q += 64) {
            const int r = q / cw, xc = q - r * cw;
            const float* g = g0 + (size_t)r * w + 4 * xc;
            const unsigned sdst = (unsigned)__cvta_generic_to_shared(s_tile + r * rw + 4 * xc);
#pragma unroll
            for (int c = t >> 6; c < kCh; c += kVarThreads / 64) {
              if (kFull || c0 + c < C)   // asynchronous copies: the whole rectangle is in flight at once
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(sdst + c * (kVarCap * 4)), "l"(g + (size_t)c * P));
              else
                *reinterpret_cast<float4*>(s_tile + c * kVarCap + r * rw + 4 * xc) = make_float4(0.f, 0.f, 0.f, 0.f);
            }
          }
        } else {
          const int area = rh * rw;
          for (int e = t & 127; e < area; e += 128) {
            const int r = e / rw, xx = e - r * rw;
            const float* g = g0 + (size_t)r * w + xx;
            const unsigned sdst = (unsigned)__cvta_generic_to_shared(s_tile + e);
#pragma unroll 4
            for (int c = t >> 7; c < kCh; c += kVarThreads / 128) {
              if (kFull || c0 + c < C)
                asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(sdst + c * (kVarCap * 4)), "l"(g + (size_t)c * P));
              else
                s_tile[c * kVarCap + e] = 0.f;
            }
          }
        }
        asm volatile("cp.async.wait_all;\n" ::: "memory");
        __syncthreads();
        if (needed) {
          // corners without weight are clamped into the rectangle: they read staged (finite) data and multiply it by 0
          const int cx0 = min(max(x0, xlo), xhi) - xo, cx1 = min(max(x0 + 1, xlo), xhi) - xo;
          const int cy0 = (min(max(y0, ylo), yhi) - ylo) * rw, cy1 = (min(max(y0 + 1, ylo), yhi) - ylo) * rw;
          const float* q_nw = s_tile + cy0 + cx0;
          const float* q_ne = s_tile + cy0 + cx1;
          const float* q_sw = s_tile + cy1 + cx0;
          const float* q_se = s_tile + cy1 + cx1;
#pragma unroll
          for (int k = 0; k < kCh; ++k) {
            float val = bl.nw * q_nw[k * kVarCap];
            val = fmaf(bl.ne, q_ne[k * kVarCap], val);
            val = fmaf(bl.sw, q_sw[k * kVarCap], val);
            val = fmaf(bl.se, q_se[k * kVarCap], val);
            sum[k] += val;
            sq[k] = fmaf(val, val, sq[k]);
          }
        }
      } else if (needed) {
        const Gather g = make_gather(ix, iy, h, w);
#pragma unroll
        for (int k = 0; k < kCh; ++k) {
          if (kFull || c0 + k < C) {
            const float val = gather_val(src + (size_t)k * P, g);
            sum[k] += val;
            sq[k] = fmaf(val, val, sq[k]);
          }
        }
      }
    }
    __syncthreads();                                        // the staged rectangle is reused by the next view
  }
  if (!active) return;
  const float inv_V = 1.0f / (float)Vt;      // x / V as x * (1 / V): within 1 ulp, far inside the 1e-4 volume tolerance
  float* out = variance + (((size_t)b * C + c0) * out_D + out_d0 + dl) * P + p;
  const size_t cstride = (size_t)out_D * P;
  const int nch = kFull ? kCh : min(kCh, C - c0);
#pragma unroll
  for (int k = 0; k < kCh; ++k, out += cstride) {
    if (kFull || k < nch) {
      const float mean = sum[k] * inv_V;
      __stcs(out, fsub(sq[k] * inv_V, fmul(mean, mean)));  // :135; streaming store: the volume is written once, never re-read here
    }
  }
}

// homo_warp, materialising (blocks/utils.py:222-268): thread = voxel (d,y,x), loops channels.
__global__ void __launch_bounds__(256) homo_warp_kernel(const float* __restrict__ src_feat, const float* __restrict__ src_proj,
                                                        const float* __restrict__ ref_proj_inv, const float* __restrict__ depths,
                                                        int C, int H, int W, int D, float* __restrict__ warped,
                                                        int32_t* __restrict__ dbg_ix0, int32_t* __restrict__ dbg_iy0) {
  __shared__ float s_M[12];
  const int b = blockIdx.y;
  const int P = H * W;
  if (threadIdx.x < 12) s_M[threadIdx.x] = matmul4_elem(src_proj + (size_t)b * 16, ref_proj_inv + (size_t)b * 16,
                                                       threadIdx.x / 4, threadIdx.x % 4);
  __syncthreads();
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (long long)D * P) return;
  const int d = (int)(idx / P), p = (int)(idx - (long long)d * P);
  const int yi = p / W, xi = p - yi * W;
  float ix, iy;
  homo_coord(s_M, xi, yi, depths[(size_t)b * D + d], H, W, ix, iy);
  int x0, y0;
  const Gather g = make_gather(ix, iy, H, W, &x0, &y0);
  if (dbg_ix0) {
    dbg_ix0[((size_t)b * D + d) * P + p] = x0;
    dbg_iy0[((size_t)b * D + d) * P + p] = y0;
  }
  const float* src = src_feat + (size_t)b * C * P;
  float* out = warped + (((size_t)b * C) * D + d) * P + p;
  for (int c = 0; c < C; ++c) out[(size_t)c * D * P] = gather_val(src + (size_t)c * P, g);
}

// ---------------------------------------------------------------------------------------------------------
// Softmax over D + expected depth + expected index + 4-bin confidence, one pass over the cost volume.
// CTA = 32 pixels x 8 plane groups: warp g streams planes g, g+8, ... for 32 adjacent pixels (128-byte rows),
// keeping the online-softmax state (m, Z, A = sum e*depth, I = sum e*index); the 8 states are merged through
// shared memory (the same merge the depth-slab partition uses across GPUs).
// ---------------------------------------------------------------------------------------------------------
constexpr int kRegWarps = 8;

struct SoftState {
  float m, Z, A, I;
};
__device__ __forceinline__ void soft_push(SoftState& s, float x, float depth, float index) {
  if (x > s.m) {
    const float r = expf(s.m - x);  // exp(-inf) = 0 on the first element
    s.Z *= r; s.A *= r; s.I *= r;
    s.m = x;
  }
  const float e = expf(x - s.m);
  s.Z += e;
  s.A = fmaf(e, depth, s.A);
  s.I = fmaf(e, index, s.I);
}
__device__ __forceinline__ void soft_merge(SoftState& a, const SoftState& b) {
  const float m = fmaxf(a.m, b.m);
  const float ra = (a.m == -INFINITY) ? 0.f : expf(a.m - m);
  const float rb = (b.m == -INFINITY) ? 0.f : expf(b.m - m);
  a.Z = a.Z * ra + b.Z * rb;
  a.A = a.A * ra + b.A * rb;
  a.I = a.I * ra + b.I * rb;
  a.m = m;
}

// mode 0: full regression (depth, confidence, index, expected index, optional prob)
// mode 1: slab partials -> partial[4,B,P]
__global__ void __launch_bounds__(kRegWarps * 32) mvs_regress_kernel(const float* __restrict__ cost, const float* __restrict__ depths,
                                                                     int B, int D, int P, int index0, int mode,
                                                                     float* __restrict__ depth_out, float* __restrict__ conf_out,
                                                                     long long* __restrict__ index_out, float* __restrict__ eidx_out,
                                                                     float* __restrict__ prob_out, float* __restrict__ partial) {
  __shared__ SoftState s_state[kRegWarps][32];
  __shared__ float s_m[32], s_Z[32];
  const int b = blockIdx.y;
  const int lane = threadIdx.x & 31, g = threadIdx.x >> 5;
  const int p = blockIdx.x * 32 + lane;
  const bool act = p < P;
  const float* x = cost + (size_t)b * D * P + p;
  const float* dv = depths + (size_t)b * D;
  SoftState st{-INFINITY, 0.f, 0.f, 0.f};
  if (act) {
    int d = g;
    constexpr int kB = 8;                                  // loads in flight per thread; one rescale per batch
    for (; d + (kB - 1) * kRegWarps < D; d += kB * kRegWarps) {
      float xv[kB], dz[kB];
#pragma unroll
      for (int k = 0; k < kB; ++k) xv[k] = __ldg(x + (size_t)(d + k * kRegWarps) * P);
#pragma unroll
      for (int k = 0; k < kB; ++k) dz[k] = __ldg(dv + d + k * kRegWarps);
      float mx = st.m;
#pragma unroll
      for (int k = 0; k < kB; ++k) mx = fmaxf(mx, xv[k]);
      const float r = __expf(st.m - mx);                   // exp(-inf) = 0 on the first batch; nan-free because mx > -inf
      st.Z *= r; st.A *= r; st.I *= r; st.m = mx;
#pragma unroll
      for (int k = 0; k < kB; ++k) {
        const float e = __expf(xv[k] - mx);
        st.Z += e;
        st.A = fmaf(e, dz[k], st.A);
        st.I = fmaf(e, (float)(index0 + d + k * kRegWarps), st.I);
      }
    }
    for (; d < D; d += kRegWarps) soft_push(st, __ldg(x + (size_t)d * P), __ldg(dv + d), (float)(index0 + d));
  }
  s_state[g][lane] = st;
  __syncthreads();
  if (g == 0) {
#pragma unroll
    for (int k = 1; k < kRegWarps; ++k) soft_merge(st, s_state[k][lane]);
    s_m[lane] = st.m;
    s_Z[lane] = st.Z;
    if (act) {
      const size_t o = (size_t)b * P + p;
      if (mode == 1) {
        const size_t BP = (size_t)B * P;
        partial[o] = st.m; partial[BP + o] = st.Z; partial[2 * BP + o] = st.A; partial[3 * BP + o] = st.I;
      } else {
        const float depth = st.A / st.Z, eidx = st.I / st.Z;
        if (depth_out) depth_out[o] = depth;
        if (eidx_out) eidx_out[o] = eidx;
        int idx = (int)eidx;  // .long() truncates (mvsnet.py:153-156)
        idx = idx < 0 ? 0 : (idx > D - 1 ? D - 1 : idx);
        if (index_out) index_out[o] = idx;
        if (conf_out) {  // prob_volume_sum4[idx] = p[idx-1] + p[idx] + p[idx+1] + p[idx+2], zero padded (:145-158)
          float acc = 0.f;
#pragma unroll
          for (int k = -1; k <= 2; ++k) {
            const int dd = idx + k;
            if (dd >= 0 && dd < D) acc += expf(__ldg(x + (size_t)dd * P) - st.m);
          }
          conf_out[o] = acc / st.Z;
        }
      }
    }
  }
  if (mode == 0 && prob_out) {
    __syncthreads();
    if (act) {
      const float m = s_m[lane], Z = s_Z[lane];
      float* po = prob_out + (size_t)b * D * P + p;
      for (int d = g; d < D; d += kRegWarps) po[(size_t)d * P] = expf(__ldg(x + (size_t)d * P) - m) / Z;
    }
  }
}

// ---------------------------------------------------------------------------------------------------------
// Vectorised form of the regression (no probability volume): CTA = 4 plane groups x 32 lanes x VEC adjacent pixels, eight
// VEC-wide loads in flight per thread; 128 threads and ~56 registers keep every CTA of a DTU-sized volume resident at once
// (one wave, no tail).  VEC = 2 needs an even pixel count (8-byte aligned rows).
// ---------------------------------------------------------------------------------------------------------
constexpr int kVecWarps = 4;

template <int VEC>
__device__ __forceinline__ void soft_accumulate(const float* __restrict__ x, const float* __restrict__ dv, int D, size_t P, int index0, int g,
                                                SoftState* st) {
  constexpr int kB = 8;
  int d = g;
  for (; d + (kB - 1) * kVecWarps < D; d += kB * kVecWarps) {
    float xv[kB][VEC], dz[kB];
#pragma unroll
    for (int k = 0; k < kB; ++k) {
      const float* q = x + (size_t)(d + k * kVecWarps) * P;
      if (VEC == 2) { const float2 t = __ldg(reinterpret_cast<const float2*>(q)); xv[k][0] = t.x; xv[k][VEC - 1] = t.y; }
      else xv[k][0] = __ldg(q);
    }
#pragma unroll
    for (int k = 0; k < kB; ++k) dz[k] = __ldg(dv + d + k * kVecWarps);
#pragma unroll
    for (int v = 0; v < VEC; ++v) {
      float mx = st[v].m;
#pragma unroll
      for (int k = 0; k < kB; ++k) mx = fmaxf(mx, xv[k][v]);
      const float r = __expf(st[v].m - mx);                // exp(-inf) = 0 on the first batch
      st[v].Z *= r; st[v].A *= r; st[v].I *= r; st[v].m = mx;
#pragma unroll
      for (int k = 0; k < kB; ++k) {
        const float e = __expf(xv[k][v] - mx);
        st[v].Z += e;
        st[v].A = fmaf(e, dz[k], st[v].A);
        st[v].I = fmaf(e, (float)(index0 + d + k * kVecWarps), st[v].I);
      }
    }
  }
  for (; d < D; d += kVecWarps)
#pragma unroll
    for (int v = 0; v < VEC; ++v) soft_push(st[v], __ldg(x + (size_t)d * P + v), __ldg(dv + d), (float)(index0 + d));
}

template <int VEC>
__global__ void __launch_bounds__(kVecWarps * 32) mvs_regress_vec_kernel(const float* __restrict__ cost, const float* __restrict__ depths, int B, int D,
                                                                         int P, int index0, int mode, float* __restrict__ depth_out,
                                                                         float* __restrict__ conf_out, long long* __restrict__ index_out,
                                                                         float* __restrict__ eidx_out, float* __restrict__ partial) {
  __shared__ SoftState s_state[kVecWarps][32 * VEC];
  const int b = blockIdx.y;
  const int lane = threadIdx.x & 31, g = threadIdx.x >> 5;
  const int p = (blockIdx.x * 32 + lane) * VEC;
  const bool act = p < P;                                    // P is a multiple of VEC: p + VEC - 1 < P too
  const float* x = cost + (size_t)b * D * P + p;
  SoftState st[VEC];
#pragma unroll
  for (int v = 0; v < VEC; ++v) st[v] = SoftState{-INFINITY, 0.f, 0.f, 0.f};
  if (act) soft_accumulate<VEC>(x, depths + (size_t)b * D, D, (size_t)P, index0, g, st);
#pragma unroll
  for (int v = 0; v < VEC; ++v) s_state[g][lane * VEC + v] = st[v];
  __syncthreads();
  // the 64 (32 * VEC) pixels of the CTA are finished by its first 32 * VEC threads, one pixel each
  const int j = threadIdx.x;
  if (j < 32 * VEC) {
    SoftState t = s_state[0][j];
#pragma unroll
    for (int k = 1; k < kVecWarps; ++k) soft_merge(t, s_state[k][j]);
    const int pp = blockIdx.x * 32 * VEC + j;
    if (pp < P) {
      const size_t o = (size_t)b * P + pp;
      if (mode == 1) {
        const size_t BP = (size_t)B * P;
        partial[o] = t.m; partial[BP + o] = t.Z; partial[2 * BP + o] = t.A; partial[3 * BP + o] = t.I;
      } else {
        const float depth = t.A / t.Z, eidx = t.I / t.Z;
        if (depth_out) depth_out[o] = depth;
        if (eidx_out) eidx_out[o] = eidx;
        int idx = (int)eidx;  // .long() truncates (mvsnet.py:153-156)
        idx = idx < 0 ? 0 : (idx > D - 1 ? D - 1 : idx);
        if (index_out) index_out[o] = idx;
        if (conf_out) {  // prob_volume_sum4[idx] = p[idx-1] + p[idx] + p[idx+1] + p[idx+2], zero padded (:145-158)
          const float* xc = cost + (size_t)b * D * P + pp;
          float acc = 0.f;
#pragma unroll
          for (int k = -1; k <= 2; ++k) {
            const int dd = idx + k;
            if (dd >= 0 && dd < D) acc += expf(__ldg(xc + (size_t)dd * P) - t.m);
          }
          conf_out[o] = acc / t.Z;
        }
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------------------
// Regression with D sharded over the G GPUs of one box, ONE kernel per rank (partition 1b).  Persistent grid (every
// CTA resident).  Phase 1: online-softmax partials (m, Z, A, I) of this rank's planes, stored into slot `rank` of EVERY rank's
// partial buffer through peer-mapped pointers.  Handshake (flags in symmetric memory, system-scope release / acquire).  Phase 2,
// every rank for all pixels: merge of the G partials, expected depth / index, and the 4-bin confidence around the global index,
// its four cost values read from whichever rank owns them (peer loads).  Second handshake: no rank leaves while another may still
// read its cost slab.  No NCCL call, no second collective for the confidence.
// ---------------------------------------------------------------------------------------------------------
struct SlabRegressArgs {
  int n, rank;
  const float* cost[RMVD_MAX_PEERS];     // [B, Dl_max, P] of every rank (slab of rank g: planes slab_lo(g) ...)
  float* parts[RMVD_MAX_PEERS];          // [n, 4, B * P] of every rank
  unsigned* flags[RMVD_MAX_PEERS];       // uint32 [2 * n] of every rank
  unsigned epoch;
  unsigned* counters;                    // local uint32 [2], zero
};

__device__ __forceinline__ void slab_handshake(const SlabRegressArgs& A, int which, unsigned nblocks) {
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned prev = atomicAdd(A.counters + which, 1u);
    if (prev == nblocks - 1) {
      A.counters[which] = 0u;
      __threadfence_system();
      for (int g = 0; g < A.n; ++g)
        asm volatile("st.release.sys.global.u32 [%0], %1;\n" ::"l"(A.flags[g] + which * A.n + A.rank), "r"(A.epoch) : "memory");
    }
    for (int g = 0; g < A.n; ++g) {
      unsigned have = 0, polls = 0;
      for (;;) {
        asm volatile("ld.acquire.sys.global.u32 %0, [%1];\n" : "=r"(have) : "l"(A.flags[A.rank] + which * A.n + g) : "memory");
        if ((int)(have - A.epoch) >= 0) break;
        __nanosleep(100);
        if (++polls > (1u << 24)) __trap();            // a peer never arrived: fail loudly rather than hang
      }
    }
  }
  __syncthreads();
}

template <int VEC>
__global__ void __launch_bounds__(kVecWarps * 32) softargmax_slab_fused_kernel(const __grid_constant__ SlabRegressArgs A, const float* __restrict__ depths,
                                                                               int B, int D, int P, float* __restrict__ depth_out,
                                                                               float* __restrict__ conf_out, long long* __restrict__ index_out,
                                                                               float* __restrict__ eidx_out) {
  __shared__ SoftState s_state[kVecWarps][32 * VEC];
  const int lane = threadIdx.x & 31, g = threadIdx.x >> 5;
  const int base = D / A.n, extra = D % A.n;
  const int lo = A.rank * base + min(A.rank, extra), Dl = base + (A.rank < extra ? 1 : 0), Dmax = base + (extra ? 1 : 0);
  const int tiles_p = (P + 32 * VEC - 1) / (32 * VEC), tiles = tiles_p * B;
  const size_t BP = (size_t)B * P;
  // phase 1
  for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
    const int b = tile / tiles_p, tp = tile - b * tiles_p;
    const int p = (tp * 32 + lane) * VEC;
    SoftState st[VEC];
#pragma unroll
    for (int v = 0; v < VEC; ++v) st[v] = SoftState{-INFINITY, 0.f, 0.f, 0.f};
    if (p < P && Dl > 0) soft_accumulate<VEC>(A.cost[A.rank] + (size_t)b * Dmax * P + p, depths + (size_t)b * D + lo, Dl, (size_t)P, lo, g, st);
    __syncthreads();
#pragma unroll
    for (int v = 0; v < VEC; ++v) s_state[g][lane * VEC + v] = st[v];
    __syncthreads();
    const int j = threadIdx.x;
    if (j < 32 * VEC) {
      SoftState t = s_state[0][j];
#pragma unroll
      for (int k = 1; k < kVecWarps; ++k) soft_merge(t, s_state[k][j]);
      const int pp = tp * 32 * VEC + j;
      if (pp < P) {
        const size_t o = (size_t)A.rank * 4 * BP + (size_t)b * P + pp;
        for (int r = 0; r < A.n; ++r) {                // this rank's partial into every rank's buffer (NVLink stores)
          float* q = A.parts[r] + o;
          q[0] = t.m; q[BP] = t.Z; q[2 * BP] = t.A; q[3 * BP] = t.I;
        }
      }
    }
  }
  slab_handshake(A, 0, gridDim.x);
  // phase 2
  const float* mine = A.parts[A.rank];
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < (long long)BP; i += (long long)gridDim.x * blockDim.x) {
    SoftState t{-INFINITY, 0.f, 0.f, 0.f};
    for (int r = 0; r < A.n; ++r) {
      const float* q = mine + (size_t)r * 4 * BP + i;
      // written by peers during this kernel: read past L1 (volatile)
      const volatile float* qv = q;
      SoftState o{qv[0], qv[BP], qv[2 * BP], qv[3 * BP]};
      soft_merge(t, o);
    }
    const float depth = t.A / t.Z, eidx = t.I / t.Z;
    if (depth_out) depth_out[i] = depth;
    if (eidx_out) eidx_out[i] = eidx;
    int idx = (int)eidx;
    idx = idx < 0 ? 0 : (idx > D - 1 ? D - 1 : idx);
    if (index_out) index_out[i] = idx;
    if (conf_out) {
      const int b = (int)(i / P), pp = (int)(i - (long long)b * P);
      float acc = 0.f;
#pragma unroll
      for (int k = -1; k <= 2; ++k) {
        const int dd = idx + k;
        if (dd >= 0 && dd < D) {
          // owner of plane dd: the first `extra` ranks hold base + 1 planes
          const int cut = extra * (base + 1);
          const int r = dd < cut ? dd / (base + 1) : extra + (dd - cut) / max(base, 1);
          const int rlo = r * base + min(r, extra);
          acc += expf(__ldg(A.cost[r] + ((size_t)b * Dmax + (dd - rlo)) * P + pp) - t.m);
        }
      }
      conf_out[i] = acc / t.Z;
    }
  }
  slab_handshake(A, 1, gridDim.x);
}

// merge G slab partials [G,4,B*P] -> merged [4,B*P] = (m, Z, depth, eidx)
__global__ void softargmax_combine_kernel(const float* __restrict__ parts, int G, long long BP, float* __restrict__ merged) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= BP) return;
  SoftState st{-INFINITY, 0.f, 0.f, 0.f};
  for (int g = 0; g < G; ++g) {
    const float* q = parts + (size_t)g * 4 * BP + i;
    SoftState o{q[0], q[BP], q[2 * BP], q[3 * BP]};
    soft_merge(st, o);
  }
  merged[i] = st.m;
  merged[BP + i] = st.Z;
  merged[2 * BP + i] = st.A / st.Z;
  merged[3 * BP + i] = st.I / st.Z;
}

// this rank's part of the 4-bin confidence around the global index
__global__ void softargmax_confidence_kernel(const float* __restrict__ cost_slab, const float* __restrict__ merged, int B,
                                             int Dl, int P, int D, int d_begin, float* __restrict__ conf_part) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  const int b = blockIdx.y;
  if (p >= P) return;
  const size_t BP = (size_t)B * P, o = (size_t)b * P + p;
  const float m = merged[o], Z = merged[BP + o], eidx = merged[3 * BP + o];
  int idx = (int)eidx;
  idx = idx < 0 ? 0 : (idx > D - 1 ? D - 1 : idx);
  float acc = 0.f;
#pragma unroll
  for (int k = -1; k <= 2; ++k) {
    const int dd = idx + k;
    if (dd >= 0 && dd < D && dd >= d_begin && dd < d_begin + Dl)
      acc += expf(cost_slab[((size_t)b * Dl + (dd - d_begin)) * P + p] - m);
  }
  conf_part[o] = acc / Z;
}

// soft_argmin + entropy (blocks/utils.py:51-68).  CTA = 32 adjacent pixels x 8 plane groups (warp g owns planes g, g + 8, ...,
// 128-byte rows, 8 loads in flight per thread).  Sweep 1: online-softmax state (m, Z, I = sum e * index) per group, merged through
// shared memory -> expected index.  Sweep 2 (the volume is L2-resident by then): probabilities, windowed probability around the
// expected index, entropy; the per-group sums meet in shared memory.
__global__ void __launch_bounds__(kRegWarps * 32) soft_argmin_kernel(const float* __restrict__ vol, int D, int P, float window,
                                                                     float* __restrict__ prob, float* __restrict__ eidx_out,
                                                                     float* __restrict__ prob_map, float* __restrict__ entropy) {
  __shared__ SoftState s_state[kRegWarps][32];
  __shared__ float s_m[32], s_Z[32], s_E[32];
  __shared__ float s_pm[kRegWarps][32], s_ent[kRegWarps][32];
  const int b = blockIdx.y;
  const int lane = threadIdx.x & 31, g = threadIdx.x >> 5;
  const int p = blockIdx.x * 32 + lane;
  const bool act = p < P;
  const float* x = vol + (size_t)b * D * P + p;
  SoftState st{-INFINITY, 0.f, 0.f, 0.f};
  constexpr int kB = 8;
  if (act) {
    int d = g;
    for (; d + (kB - 1) * kRegWarps < D; d += kB * kRegWarps) {
      float xv[kB];
#pragma unroll
      for (int k = 0; k < kB; ++k) xv[k] = __ldg(x + (size_t)(d + k * kRegWarps) * P);
      float mx = st.m;
#pragma unroll
      for (int k = 0; k < kB; ++k) mx = fmaxf(mx, xv[k]);
      const float r = expf(st.m - mx);                     // exp(-inf) = 0 on the first batch
      st.Z *= r; st.I *= r; st.m = mx;
#pragma unroll
      for (int k = 0; k < kB; ++k) {
        const float e = expf(xv[k] - mx);
        st.Z += e;
        st.I = fmaf(e, (float)(d + k * kRegWarps), st.I);
      }
    }
    for (; d < D; d += kRegWarps) soft_push(st, __ldg(x + (size_t)d * P), 0.f, (float)d);
  }
  s_state[g][lane] = st;
  __syncthreads();
  if (g == 0) {
#pragma unroll
    for (int k = 1; k < kRegWarps; ++k) soft_merge(st, s_state[k][lane]);
    const float E = st.I / st.Z;
    s_m[lane] = st.m; s_Z[lane] = st.Z; s_E[lane] = E;
    if (act) eidx_out[(size_t)b * P + p] = E;
  }
  if (!prob && !prob_map && !entropy) return;
  __syncthreads();
  float pm = 0.f, ent = 0.f;
  if (act) {
    const float m = s_m[lane], inv_Z = 1.0f / s_Z[lane], E = s_E[lane];
    float* po = prob ? prob + (size_t)b * D * P + p : nullptr;
    int d = g;
    for (; d + (kB - 1) * kRegWarps < D; d += kB * kRegWarps) {
      float xv[kB];
#pragma unroll
      for (int k = 0; k < kB; ++k) xv[k] = __ldg(x + (size_t)(d + k * kRegWarps) * P);
#pragma unroll
      for (int k = 0; k < kB; ++k) {
        const int dd = d + k * kRegWarps;
        const float pd = expf(xv[k] - m) * inv_Z;
        if (po) __stcs(po + (size_t)dd * P, pd);
        if (fabsf((float)dd - E) <= window) pm += pd;
        ent = fmaf(-pd, logf(fminf(fmaxf(pd, 1e-9f), 1.0f)), ent);
      }
    }
    for (; d < D; d += kRegWarps) {
      const float pd = expf(__ldg(x + (size_t)d * P) - m) * inv_Z;
      if (po) __stcs(po + (size_t)d * P, pd);
      if (fabsf((float)d - E) <= window) pm += pd;
      ent = fmaf(-pd, logf(fminf(fmaxf(pd, 1e-9f), 1.0f)), ent);
    }
  }
  if (!prob_map && !entropy) return;
  s_pm[g][lane] = pm; s_ent[g][lane] = ent;
  __syncthreads();
  if (g == 0 && act) {
#pragma unroll
    for (int k = 1; k < kRegWarps; ++k) { pm += s_pm[k][lane]; ent += s_ent[k][lane]; }
    const size_t o = (size_t)b * P + p;
    if (prob_map) prob_map[o] = pm;
    if (entropy) entropy[o] = ent;
  }
}

// ---------------------------------------------------------------------------------------------------------
// Plain reductions over D of a [B,D,h,w] volume, no softmax: depth_regression (blocks/utils.py:271-274: sum_d p * depth, depth
// values per plane [B,D] or per pixel [B,D,h,w] as cvp_mvsnet.py:168 passes them) and entropy (:67-68:
// sum_d -v * log(clamp(v, 1e-9, 1))).  Thread = pixel, lanes = adjacent pixels (128-byte rows), 4 planes in flight.
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) volume_reduce_kernel(const float* __restrict__ vol, const float* __restrict__ values, int D, int P,
                                                            int mode, float* __restrict__ out) {
  const int b = blockIdx.y;
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= P) return;
  const float* v = vol + (size_t)b * D * P + p;
  float acc = 0.f;
  if (mode == 2) {                                   // entropy
    for (int d = 0; d < D; ++d) {
      const float x = __ldg(v + (size_t)d * P);
      acc = fadd(acc, fmul(-x, logf(fminf(fmaxf(x, 1e-9f), 1.0f))));
    }
  } else if (mode == 1) {                            // per-pixel values [B,D,h,w]
    const float* q = values + (size_t)b * D * P + p;
    for (int d = 0; d < D; ++d) acc = fadd(acc, fmul(__ldg(v + (size_t)d * P), __ldg(q + (size_t)d * P)));
  } else {                                           // per-plane values [B,D]
    const float* q = values + (size_t)b * D;
    for (int d = 0; d < D; ++d) acc = fadd(acc, fmul(__ldg(v + (size_t)d * P), __ldg(q + d)));
  }
  out[(size_t)b * P + p] = acc;
}

}  // namespace rmvd

using namespace rmvd;

extern "C" int rmvd_mvs_variance_fwd(const float* feats, const float* proj, const float* depths, int B, int Vt, int C,
                                     int h, int w, int D, int d_begin, int d_end, float* variance, int out_D,
                                     int out_d0, void* stream) {
  if (!feats || !proj || !depths || !variance) return RMVD_E_NULL;
  if (B <= 0 || Vt <= 0 || C <= 0 || h <= 0 || w <= 0 || D <= 0 || out_D <= 0 || B > 65535) return RMVD_E_SHAPE;
  if (Vt - 1 > kMaxMvsViews) return RMVD_E_SHAPE;
  if (d_begin < 0 || d_end > D || d_begin > d_end || out_d0 < 0 || out_d0 + (d_end - d_begin) > out_D)
    return RMVD_E_RANGE;
  if (d_begin == d_end) return RMVD_OK;
  static const bool direct = getenv("RMVD_MVS_VARIANCE_DIRECT") != nullptr;   // A/B switch: the un-staged kernel
  if (direct) {
    const long long tot = (long long)(d_end - d_begin) * h * w;
    constexpr int kCh = 8;
    dim3 grid((unsigned)((tot + 255) / 256), (C + kCh - 1) / kCh, B);
    mvs_variance_kernel<kCh><<<grid, 256, 0, (cudaStream_t)stream>>>(feats, proj, depths, B, Vt, C, h, w, D, d_begin,
                                                                      d_end, variance, out_D, out_d0);
    return (int)cudaGetLastError();
  }
  const int tiles_x = (w + kVarPx - 1) / kVarPx;
  const int ndb = (d_end - d_begin + kVarDB - 1) / kVarDB;
  auto launch = [&](auto kernel, int kCh) -> int {
    const int ncg = (C + kCh - 1) / kCh;
    if ((long long)ndb * ncg > 65535) return RMVD_E_SHAPE;
    const size_t smem = sizeof(float) * kCh * kVarCap;
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    dim3 grid((unsigned)(tiles_x * h), (unsigned)(ndb * ncg), B);
    const int vec16 = (w % 4 == 0) && ((uintptr_t)feats % 16 == 0);
    kernel<<<grid, kVarThreads, smem, (cudaStream_t)stream>>>(feats, proj, depths, B, Vt, C, h, w, D, d_begin, d_end, variance,
                                                              out_D, out_d0, tiles_x, ncg, vec16);
    return (int)cudaGetLastError();
  };
  if (C > 16) return (C % 32 == 0) ? launch(mvs_variance_tile_kernel<32, true>, 32) : launch(mvs_variance_tile_kernel<32, false>, 32);
  return (C % 16 == 0) ? launch(mvs_variance_tile_kernel<16, true>, 16) : launch(mvs_variance_tile_kernel<16, false>, 16);
}

extern "C" int rmvd_homo_warp_fwd(const float* src_feat, const float* src_proj, const float* ref_proj_inv,
                                  const float* depths, int B, int C, int H, int W, int D, float* warped,
                                  int32_t* dbg_ix0, int32_t* dbg_iy0, void* stream) {
  if (!src_feat || !src_proj || !ref_proj_inv || !depths || !warped) return RMVD_E_NULL;
  if ((dbg_ix0 == nullptr) != (dbg_iy0 == nullptr)) return RMVD_E_NULL;
  if (B <= 0 || C <= 0 || H <= 0 || W <= 0 || D <= 0 || B > 65535) return RMVD_E_SHAPE;
  const long long tot = (long long)D * H * W;
  dim3 grid((unsigned)((tot + 255) / 256), B);
  homo_warp_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(src_feat, src_proj, ref_proj_inv, depths, C, H, W, D, warped,
                                                          dbg_ix0, dbg_iy0);
  return (int)cudaGetLastError();
}

extern "C" int rmvd_mvs_regress_fwd(const float* cost, const float* depths, int B, int D, int h, int w, float* depth,
                                    float* confidence, int64_t* depth_index, float* expected_index, float* prob,
                                    void* stream) {
  if (!cost || !depths) return RMVD_E_NULL;
  if (B <= 0 || D <= 0 || h <= 0 || w <= 0 || B > 65535) return RMVD_E_SHAPE;
  const int P = h * w;
  if (!prob) {
    if (P % 2 == 0) {
      dim3 grid((P + 63) / 64, B);
      mvs_regress_vec_kernel<2><<<grid, kVecWarps * 32, 0, (cudaStream_t)stream>>>(cost, depths, B, D, P, 0, 0, depth, confidence,
                                                                                   (long long*)depth_index, expected_index, nullptr);
    } else {
      dim3 grid((P + 31) / 32, B);
      mvs_regress_vec_kernel<1><<<grid, kVecWarps * 32, 0, (cudaStream_t)stream>>>(cost, depths, B, D, P, 0, 0, depth, confidence,
                                                                                   (long long*)depth_index, expected_index, nullptr);
    }
    return (int)cudaGetLastError();
  }
  dim3 grid((P + 31) / 32, B);
  mvs_regress_kernel<<<grid, kRegWarps * 32, 0, (cudaStream_t)stream>>>(cost, depths, B, D, P, 0, 0, depth, confidence,
                                                                        (long long*)depth_index, expected_index, prob,
                                                                        nullptr);
  return (int)cudaGetLastError();
}

extern "C" int rmvd_softargmax_partial(const float* cost_slab, const float* depths_slab, int B, int Dl, int h, int w,
                                       int d_begin, float* partial, void* stream) {
  if (!cost_slab || !depths_slab || !partial) return RMVD_E_NULL;
  if (B <= 0 || Dl <= 0 || h <= 0 || w <= 0 || d_begin < 0 || B > 65535) return RMVD_E_SHAPE;
  const int P = h * w;
  if (P % 2 == 0) {
    dim3 grid((P + 63) / 64, B);
    mvs_regress_vec_kernel<2><<<grid, kVecWarps * 32, 0, (cudaStream_t)stream>>>(cost_slab, depths_slab, B, Dl, P, d_begin, 1, nullptr, nullptr,
                                                                                 nullptr, nullptr, partial);
  } else {
    dim3 grid((P + 31) / 32, B);
    mvs_regress_vec_kernel<1><<<grid, kVecWarps * 32, 0, (cudaStream_t)stream>>>(cost_slab, depths_slab, B, Dl, P, d_begin, 1, nullptr, nullptr,
                                                                                 nullptr, nullptr, partial);
  }
  return (int)cudaGetLastError();
}

extern "C" int rmvd_softargmax_combine(const float* partials, int G, int B, int h, int w, float* merged, void* stream) {
  if (!partials || !merged) return RMVD_E_NULL;
  if (G <= 0 || B <= 0 || h <= 0 || w <= 0) return RMVD_E_SHAPE;
  const long long BP = (long long)B * h * w;
  softargmax_combine_kernel<<<(unsigned)((BP + 255) / 256), 256, 0, (cudaStream_t)stream>>>(partials, G, BP, merged);
  return (int)cudaGetLastError();
}

extern "C" int rmvd_softargmax_confidence(const float* cost_slab, const float* merged, int B, int Dl, int h, int w,
                                          int D, int d_begin, float* conf_part, void* stream) {
  if (!cost_slab || !merged || !conf_part) return RMVD_E_NULL;
  if (B <= 0 || Dl <= 0 || h <= 0 || w <= 0 || D <= 0 || d_begin < 0 || B > 65535) return RMVD_E_SHAPE;
  const int P = h * w;
  dim3 grid((P + 255) / 256, B);
  softargmax_confidence_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(cost_slab, merged, B, Dl, P, D, d_begin, conf_part);
  return (int)cudaGetLastError();
}

extern "C" int rmvd_softargmax_slab_fused(const float* const* cost_peers, const float* depths, int B, int D, int h, int w, int num_peers,
                                          int rank, float* const* parts_peers, uint32_t* const* flag_peers, uint32_t epoch,
                                          uint32_t* local_counters, float* depth, float* confidence, int64_t* depth_index,
                                          float* expected_index, void* stream) {
  if (!cost_peers || !depths || !parts_peers || !flag_peers || !local_counters) return RMVD_E_NULL;
  if (B <= 0 || D <= 0 || h <= 0 || w <= 0) return RMVD_E_SHAPE;
  if (num_peers < 1 || num_peers > RMVD_MAX_PEERS || rank < 0 || rank >= num_peers) return RMVD_E_RANGE;
  SlabRegressArgs A{};
  A.n = num_peers; A.rank = rank; A.epoch = epoch; A.counters = local_counters;
  for (int g = 0; g < num_peers; ++g) {
    if (!cost_peers[g] || !parts_peers[g] || !flag_peers[g]) return RMVD_E_NULL;
    A.cost[g] = cost_peers[g]; A.parts[g] = parts_peers[g]; A.flags[g] = flag_peers[g];
  }
  const int P = h * w;
  static int slots[2] = {0, 0};                        // resident CTAs of the two instantiations on this device
  const int vi = (P % 2 == 0) ? 1 : 0;
  if (!slots[vi]) {
    int dev = 0, sms = 0, per_sm = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (e == cudaSuccess)
      e = vi ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, softargmax_slab_fused_kernel<2>, kVecWarps * 32, 0)
             : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, softargmax_slab_fused_kernel<1>, kVecWarps * 32, 0);
    if (e != cudaSuccess) return (int)e;
    slots[vi] = sms * per_sm;
  }
  const int VEC = vi ? 2 : 1;
  const long long tiles = (long long)((P + 32 * VEC - 1) / (32 * VEC)) * B;
  // every CTA waits inside the kernel for the other ranks: the grid must be resident as a whole (leave a quarter of the slots free)
  const int grid = (int)(tiles < (long long)(slots[vi] * 3 / 4) ? tiles : (long long)(slots[vi] * 3 / 4));
  if (vi)
    softargmax_slab_fused_kernel<2><<<grid, kVecWarps * 32, 0, (cudaStream_t)stream>>>(A, depths, B, D, P, depth, confidence, (long long*)depth_index,
                                                                                       expected_index);
  else
    softargmax_slab_fused_kernel<1><<<grid, kVecWarps * 32, 0, (cudaStream_t)stream>>>(A, depths, B, D, P, depth, confidence, (long long*)depth_index,
                                                                                       expected_index);
  return (int)cudaGetLastError();
}

extern "C" int rmvd_soft_argmin_fwd(const float* volume, int B, int D, int h, int w, float window, float* prob,
                                    float* expected_index, float* prob_map, float* entropy, void* stream) {
  if (!volume || !expected_index) return RMVD_E_NULL;
  if (B <= 0 || D <= 0 || h <= 0 || w <= 0 || B > 65535) return RMVD_E_SHAPE;
  const int P = h * w;
  dim3 grid((P + 31) / 32, B);
  soft_argmin_kernel<<<grid, kRegWarps * 32, 0, (cudaStream_t)stream>>>(volume, D, P, window, prob, expected_index, prob_map, entropy);
  return (int)cudaGetLastError();
}

extern "C" int rmvd_abi_version(void) { return RMVD_ABI_VERSION; }

extern "C" const char* rmvd_status_string(int status) {
  switch (status) {
    case RMVD_OK: return "ok";
    case RMVD_E_NULL: return "required pointer is NULL";
    case RMVD_E_SHAPE: return "non-positive or inconsistent size";
    case RMVD_E_MODE: return "unknown mode";
    case RMVD_E_WORKSPACE: return "workspace missing or too small";
    case RMVD_E_RANGE: return "plane range outside the schedule";
    case RMVD_E_STATIC_LIMIT: return "problem needs more scratch than the library's static scratch array (RMVD_STATIC_SCRATCH_BYTES)";
    default: return status > 0 ? cudaGetErrorString((cudaError_t)status) : "unknown status";
  }
}

// CVP-MVSNet projection set-up (cvp_mvsnet_components.py:398-405 / :204-211) on the device: thread = (sample, view).
// View 0 gets inverse(ref_proj), views v >= 1 get src_proj, each a 4x4 with the row (0, 0, 0, 1): what
// rmvd_cvp_proj_cost_fwd / rmvd_homo_warp_fwd take.  The affine inverse is evaluated in double and rounded once.
__global__ void cvp_projection_kernel(const float* __restrict__ ref_in, const float* __restrict__ src_in,
                                      const float* __restrict__ ref_ex, const float* __restrict__ src_ex, int B, int nsrc,
                                      float* __restrict__ proj) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * (nsrc + 1)) return;
  const int b = i / (nsrc + 1), v = i % (nsrc + 1);
  const float* K = (v == 0) ? ref_in + (size_t)b * 9 : src_in + ((size_t)b * nsrc + v - 1) * 9;
  const float* E = (v == 0) ? ref_ex + (size_t)b * 16 : src_ex + ((size_t)b * nsrc + v - 1) * 16;
  float M[12];  // K @ E[:3, :]
#pragma unroll
  for (int r = 0; r < 3; ++r)
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      float acc = fmul(K[r * 3 + 0], E[0 * 4 + c]);
      acc = __fmaf_rn(K[r * 3 + 1], E[1 * 4 + c], acc);
      M[r * 4 + c] = __fmaf_rn(K[r * 3 + 2], E[2 * 4 + c], acc);
    }
  float* out = proj + (size_t)i * 16;
  if (v > 0) {
#pragma unroll
    for (int k = 0; k < 12; ++k) out[k] = M[k];
  } else {  // inverse of [[A, t], [0, 1]] = [[A^-1, -A^-1 t], [0, 1]]
    const double a = M[0], bb = M[1], c = M[2], d = M[4], e = M[5], f = M[6], g = M[8], h = M[9], k = M[10];
    const double c00 = e * k - f * h, c01 = -(d * k - f * g), c02 = d * h - e * g;
    const double det = a * c00 + bb * c01 + c * c02;
    const double inv[9] = {c00 / det, -(bb * k - c * h) / det, (bb * f - c * e) / det,
                           c01 / det, (a * k - c * g) / det, -(a * f - c * d) / det,
                           c02 / det, -(a * h - bb * g) / det, (a * e - bb * d) / det};
    const double t[3] = {M[3], M[7], M[11]};
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      out[r * 4 + 0] = (float)inv[r * 3 + 0];
      out[r * 4 + 1] = (float)inv[r * 3 + 1];
      out[r * 4 + 2] = (float)inv[r * 3 + 2];
      out[r * 4 + 3] = (float)(-(inv[r * 3 + 0] * t[0] + inv[r * 3 + 1] * t[1] + inv[r * 3 + 2] * t[2]));
    }
  }
  out[12] = 0.f; out[13] = 0.f; out[14] = 0.f; out[15] = 1.f;
}

extern "C" int rmvd_cvp_projections(const float* ref_in, const float* src_in, const float* ref_ex, const float* src_ex, int B,
                                    int nsrc, float* proj, void* stream) {
  if (!ref_in || !src_in || !ref_ex || !src_ex || !proj) return RMVD_E_NULL;
  if (B <= 0 || nsrc <= 0) return RMVD_E_SHAPE;
  const int tot = B * (nsrc + 1);
  cvp_projection_kernel<<<(tot + 127) / 128, 128, 0, (cudaStream_t)stream>>>(ref_in, src_in, ref_ex, src_ex, B, nsrc, proj);
  return (int)cudaGetLastError();
}

extern "C" int rmvd_cvp_proj_cost_fwd(const float* feats, const float* proj, const float* depth_hypos, int B, int Vt, int C,
                                      int h, int w, int D, float* cost, void* stream) {
  if (!feats || !proj || !depth_hypos || !cost) return RMVD_E_NULL;
  if (B <= 0 || Vt <= 0 || C <= 0 || h <= 0 || w <= 0 || D <= 0 || B > 65535) return RMVD_E_SHAPE;
  if (Vt - 1 > kMaxMvsViews) return RMVD_E_SHAPE;
  const long long tot = (long long)D * h * w;
  constexpr int kCh = 8;
  dim3 grid((unsigned)((tot + 255) / 256), (C + kCh - 1) / kCh, B);
  mvs_variance_kernel<kCh, true><<<grid, 256, 0, (cudaStream_t)stream>>>(feats, proj, depth_hypos, B, Vt, C, h, w, D, 0, D,
                                                                            cost, D, 0);
  return (int)cudaGetLastError();
}

extern "C" int rmvd_volume_reduce_fwd(const float* volume, const float* values, int B, int D, int h, int w, int mode, float* out,
                                      void* stream) {
  if (!volume || !out || (mode != RMVD_REDUCE_ENTROPY && !values)) return RMVD_E_NULL;
  if (B <= 0 || D <= 0 || h <= 0 || w <= 0 || B > 65535) return RMVD_E_SHAPE;
  if (mode < 0 || mode > 2) return RMVD_E_MODE;
  const int P = h * w;
  dim3 grid((P + 255) / 256, B);
  volume_reduce_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(volume, values, D, P, mode, out);
  return (int)cudaGetLastError();
}
