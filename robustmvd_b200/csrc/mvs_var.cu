// MVSNet variance volume, channels-last staged forms (sm_100a; the warp-specialised kernel at the end of the namespace is the
// default): homo_warp (blocks/utils.py:222-268) fused with the volume assembly of mvsnet.py:123-136 for C a multiple of 32.  What changed against mvs_variance_tile_kernel (mvs.cu), which spent
// its issue slots on 128 scalar shared-memory loads per (voxel, view):
//
//   * the source maps are repacked once per call into channels-last [B, Vt-1, h*w, C] (nchw_to_cl_kernel, 2 % of the bytes the
//     volume writes), so a texel's 32 channels are one 128-byte line: the staged rectangle is copied with fully coalesced
//     16-byte cp.async (zero-filled outside the map = grid_sample's zeros padding, no per-corner bounds tests afterwards), and
//     a corner costs 8 LDS.128 instead of 32 LDS.32;
//   * shared-memory layout "chunk-major": [chunk of 4 channels][texel] 16-byte cells with a chunk pitch of kCap * 16 bytes,
//     kCap = 1 (mod 8): the copy's quarter-warps (8 chunks of one texel) and the gather's quarter-warps (one chunk of 8
//     neighbouring texels) both cover all 32 banks, and the chunk offset is an immediate of the load (no address arithmetic in
//     the channel loop);
//   * packed fp32 arithmetic (FFMA2 / FMUL2 / FADD2, two channels per instruction) for the interpolation and the running sums;
//   * the rectangle of a step (plane group, view) is planned from the tile's corner pixels and copied ahead of the gather: by every
//     warp one step ahead with one CTA barrier per step (mvs_variance_cl_kernel), or by a producer warpgroup through a ring of four
//     buffers on full / empty mbarriers (mvs_variance_ws_kernel, the default): no exposed copy latency;
//   * tiles of 16 x 4 key pixels (staged texels per voxel: 0.19-0.24 at the DTU shape against 0.34-0.51 for 64 x 1).
//
// Thread = one voxel (key pixel, plane) x 32 channels: sum and sum of squares of the source views (mvsnet.py:124-133) in 64 registers
// across the views, the key view's terms added when a plane group is written; the sampling coordinate comes from the reference's own
// chain with shared reciprocals (homo_coord_fast).  CTA = 64 key pixels x kPL planes, loops over plane groups and views.  A voxel whose
// cell is not inside the planned rectangle (strong rotation between the views, irregular geometry) is gathered from the channels-last
// copy in global memory, with bounds tests.
#include <climits>

#include "mvs_coord.cuh"

namespace rmvd {
namespace varcl {

constexpr int kTW = 16, kTH = 4, kPx = kTW * kTH;   // key pixels per CTA
constexpr int kCG = 32;                             // channels per CTA
constexpr int kChunks = kCG / 4;                    // 16-byte chunks per texel
constexpr int kCap = 321;                           // staged texels per buffer; = 1 (mod 8), see above
constexpr int kChunkPitch = kCap * 16;              // bytes between the chunk planes of a buffer
constexpr int kBufBytes = kChunks * kChunkPitch;    // 41 088 bytes
constexpr int kNone = INT_MAX;

struct Loc {            // sampling footprint of one (voxel, view)
  int x0, y0;           // integer corner (floor of the un-normalised grid_sample coordinate)
  float wx, wy;         // fractional parts
  bool needed;          // some corner lies inside the source map
};
struct Rect {
  int rx, ry, rw;
  bool fits;
};

__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }

// NCHW -> channels-last for the source views (views 1..Vt-1) of every sample in one launch.  grid = (ceil(P / 32), C / 32, B * (Vt - 1)).
__global__ void __launch_bounds__(256) nchw_to_cl_kernel(const float* __restrict__ feats, float* __restrict__ dst, int Vt, int C, int P) {
  __shared__ float tile[32][33];
  const int z = blockIdx.z, nV = Vt - 1;
  const int b = z / nV, v = 1 + z - b * nV;
  const float* src = feats + (((size_t)b * Vt + v) * C + blockIdx.y * 32) * P;
  float* out = dst + ((size_t)z * P) * C + blockIdx.y * 32;
  const int p0 = blockIdx.x * 32, tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
#pragma unroll
  for (int k = ty; k < 32; k += 8) tile[k][tx] = (p0 + tx < P) ? __ldg(src + (size_t)k * P + p0 + tx) : 0.f;
  __syncthreads();
#pragma unroll
  for (int k = ty; k < 32; k += 8)
    if (p0 + k < P) out[(size_t)(p0 + k) * C + tx] = tile[tx][k];
}

// 4 channels of one voxel: val = nw * A + ne * B + sw * C + se * D (bilinear grid_sample), sum += val, sq += val^2 (mvsnet.py:128-133)
template <bool kPacked>
__device__ __forceinline__ void accumulate4(const float4& A, const float4& B, const float4& Cq, const float4& Dq, float nw, float ne, float sw,
                                            float se, float2* sum, float2* sq) {
  if (kPacked) {
    const float2 nw2 = make_float2(nw, nw), ne2 = make_float2(ne, ne), sw2 = make_float2(sw, sw), se2 = make_float2(se, se);
    float2 lo = __fmul2_rn(nw2, make_float2(A.x, A.y)), hi = __fmul2_rn(nw2, make_float2(A.z, A.w));
    lo = __ffma2_rn(ne2, make_float2(B.x, B.y), lo);   hi = __ffma2_rn(ne2, make_float2(B.z, B.w), hi);
    lo = __ffma2_rn(sw2, make_float2(Cq.x, Cq.y), lo); hi = __ffma2_rn(sw2, make_float2(Cq.z, Cq.w), hi);
    lo = __ffma2_rn(se2, make_float2(Dq.x, Dq.y), lo); hi = __ffma2_rn(se2, make_float2(Dq.z, Dq.w), hi);
    sum[0] = __fadd2_rn(sum[0], lo);     sum[1] = __fadd2_rn(sum[1], hi);
    sq[0] = __ffma2_rn(lo, lo, sq[0]);   sq[1] = __ffma2_rn(hi, hi, sq[1]);
  } else {
    const float a[4] = {A.x, A.y, A.z, A.w}, bb[4] = {B.x, B.y, B.z, B.w}, c[4] = {Cq.x, Cq.y, Cq.z, Cq.w}, d[4] = {Dq.x, Dq.y, Dq.z, Dq.w};
    float* s = reinterpret_cast<float*>(sum);
    float* q = reinterpret_cast<float*>(sq);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      float val = nw * a[k];
      val = fmaf(ne, bb[k], val);
      val = fmaf(sw, c[k], val);
      val = fmaf(se, d[k], val);
      s[k] += val;
      q[k] = fmaf(val, val, q[k]);
    }
  }
}

// Sampling coordinate with the reference's chain (homo_coord, blocks/utils.py:240-264), divisions through shared reciprocals without
// the range guards of homo_coord_shared: bit-identical to div.rn while every intermediate stays inside [2^-60, 2^60], which image-scale
// geometry does; beyond that range the sample is far outside the map or the point is at the camera centre, and the variance is a
// continuous function of the coordinate in any case (unlike an index or a mask, it has nothing to flip).
__device__ __forceinline__ void homo_coord_fast(const float* M, float xf, float yf, float d, int H, int W, float y_w, float y_h, float& ix,
                                                float& iy) {
  const float gx = fmul(xf, d), gy = fmul(yf, d);
  float p[3];
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    float acc = fmul(M[r * 4 + 0], gx);
    acc = __fmaf_rn(M[r * 4 + 1], gy, acc);
    acc = __fmaf_rn(M[r * 4 + 2], d, acc);
    p[r] = fadd(acc, M[r * 4 + 3]);
  }
  const float y = rcp_refined(p[2]);
  const float sx = div_shared(p[0], p[2], y), sy = div_shared(p[1], p[2], y);
  const float xn = fsub(div_shared(sx, (float)(W - 1) * 0.5f, y_w), 1.0f);
  const float yn = fsub(div_shared(sy, (float)(H - 1) * 0.5f, y_h), 1.0f);
  ix = grid_unnormalize(xn, W);
  iy = grid_unnormalize(yn, H);
}

// Pipeline of one CTA (tile of 64 key pixels, a range of plane groups), step = (plane group, source view):
//   during step s   warp 0 plans the rectangle of step s + 2: the sampling coordinates of the tile's 4 corner pixels at each of the
//                   group's planes (exact chain) bound every sample of the step, because at a fixed plane the coordinate is a
//                   projective function of the pixel (extremes at the corners while the denominator keeps its sign);
//                   the first threads issue the copies of step s + 1's rectangle (one texel per thread: 8 x cp.async of 16 bytes);
//                   every thread locates its voxel for step s and gathers it from step s's rectangle;
//   one barrier     (copies of s + 1 landed, plan of s + 2 visible, everyone done with s's buffer).
// A voxel whose corner cell is not inside the planned rectangle (rectangle too large, irregular geometry, rounding at the rim) is
// gathered from the channels-last copy in global memory with bounds tests: the plan is a performance hint, never a correctness input.
template <int kPL, bool kPacked>
__global__ void __launch_bounds__(kPx * kPL, (kPL <= 4) ? 2 : 1) mvs_variance_cl_kernel(
    const float* __restrict__ feats, const float* __restrict__ src_cl, const float* __restrict__ proj, const float* __restrict__ depths,
    int B, int Vt, int C, int h, int w, int D, int d_begin, int d_end, float* __restrict__ variance, int out_D, int out_d0, int tiles_x,
    int ncg, int nsplit) {
  constexpr int kThreads = kPx * kPL;
  static_assert(4 * kPL <= 32, "one warp plans a step: 4 corner pixels x kPL planes");
  extern __shared__ __align__(128) unsigned char s_buf[];   // two staged rectangles
  __shared__ float s_M[kMaxMvsViews][12];
  __shared__ __align__(16) int s_rect[3][4];   // {rx, ry, rw, rh} of a step, three sets in rotation; rw = 0: nothing staged
  __shared__ float2 s_ref[(kCG / 2) * kPx];   // key-view features of the tile, [channel pair][tile pixel]
  const int t = threadIdx.x, lane = t & 31;
  const int b = blockIdx.z;
  const int P = h * w, nV = Vt - 1;
  const int tyi = blockIdx.x / tiles_x, txi = blockIdx.x - tyi * tiles_x;
  const int split = blockIdx.y / ncg, cg = blockIdx.y - split * ncg;
  const int c0 = cg * kCG;
  for (int e = t; e < nV * 12; e += kThreads) {
    const int v = 1 + e / 12, i = (e % 12) / 4, j = e % 4;
    s_M[v - 1][i * 4 + j] = matmul4_elem(proj + ((size_t)b * Vt + v) * 16, proj + ((size_t)b * Vt) * 16, i, j);
  }
  __syncthreads();

  const int px = t & (kPx - 1), pl = t / kPx;
  const int xi = txi * kTW + (px % kTW), yi = tyi * kTH + (px / kTW);
  const bool in_img = (xi < w) && (yi < h);
  const int p = min(yi, h - 1) * w + min(xi, w - 1);
  const int Dl = d_end - d_begin;
  const int npg = (Dl + kPL - 1) / kPL, pg_per = (npg + nsplit - 1) / nsplit;
  const int pg0 = split * pg_per, pg1 = min(npg, pg0 + pg_per);
  if (pg0 >= pg1) return;                                   // uniform
  for (int e = t; e < kCG * kPx; e += kThreads) {           // visible after the first barrier below
    const int k = e / kPx, q = e - k * kPx;
    const int qx = txi * kTW + (q % kTW), qy = tyi * kTH + (q / kTW);
    reinterpret_cast<float*>(s_ref)[((k >> 1) * kPx + q) * 2 + (k & 1)] =
        (qx < w && qy < h) ? __ldg(feats + (((size_t)b * Vt) * C + c0 + k) * P + (size_t)qy * w + qx) : 0.f;
  }
  const int nsteps = (pg1 - pg0) * nV;
  const float y_w = rcp_refined((float)(w - 1) * 0.5f), y_h = rcp_refined((float)(h - 1) * 0.5f);
  const float wm = (float)(w - 1), hm = (float)(h - 1);
  const float xf = (float)xi, yf = (float)yi;
  const unsigned sbase = smem_u32(s_buf);
  const float* src_b = src_cl + ((size_t)b * nV * P) * C + c0;
  const float* dep_b = depths + (size_t)b * D + d_begin;

  // warp 0: rectangle of (plane group pg, view v) -> s_rect[slot]
  auto plan = [&](int pg, int v, int slot) {
    const int corner = lane & 3, cpl = lane >> 2;
    const int dl = pg * kPL + cpl;
    const bool valid = (cpl < kPL) && (dl < Dl);
    const int cx = (corner & 1) ? min(txi * kTW + kTW - 1, w - 1) : txi * kTW;
    const int cy = (corner & 2) ? min(tyi * kTH + kTH - 1, h - 1) : tyi * kTH;
    const float d = __ldg(dep_b + min(dl, Dl - 1));
    const float* M = s_M[v - 1];
    float ix, iy;
    homo_coord_fast(M, (float)cx, (float)cy, d, h, w, y_w, y_h, ix, iy);
    const float den = fadd(__fmaf_rn(M[10], d, __fmaf_rn(M[9], fmul((float)cy, d), fmul(M[8], fmul((float)cx, d)))), M[11]);
    const bool pos = den > 0.f;
    const bool pos0 = __shfl_sync(kFullMask, pos, lane & ~3);
    const bool regular = !valid || (isfinite(ix) && isfinite(iy) && (den != 0.f) && (pos == pos0));
    const bool all_regular = __all_sync(kFullMask, regular);
    // clamp before the conversion; margin 2^-8 px for the rounding differences between a corner's chain and an interior pixel's
    const float lo_x = fminf(fmaxf(floorf(ix - 0.00390625f), -1.f), (float)w), hi_x = fminf(fmaxf(floorf(ix + 0.00390625f) + 1.f, -1.f), (float)w);
    const float lo_y = fminf(fmaxf(floorf(iy - 0.00390625f), -1.f), (float)h), hi_y = fminf(fmaxf(floorf(iy + 0.00390625f) + 1.f, -1.f), (float)h);
    const bool use = valid && all_regular;
    const int xlo = __reduce_min_sync(kFullMask, use ? (int)lo_x : kNone), xhi = __reduce_max_sync(kFullMask, use ? (int)hi_x : -kNone);
    const int ylo = __reduce_min_sync(kFullMask, use ? (int)lo_y : kNone), yhi = __reduce_max_sync(kFullMask, use ? (int)hi_y : -kNone);
    if (lane == 0) {
      int rw = 0, rh = 0;
      // some cell with a corner inside the map: x0 in [-1, w-1] <=> columns [x0, x0 + 1] meet [0, w-1]
      if (all_regular && xlo != kNone && xlo <= w - 1 && xhi >= 0 && ylo <= h - 1 && yhi >= 0) {
        rw = xhi - xlo + 1;
        rh = yhi - ylo + 1;
        if (rw * rh > kCap) rw = rh = 0;
      }
      s_rect[slot][0] = xlo; s_rect[slot][1] = ylo; s_rect[slot][2] = rw; s_rect[slot][3] = rh;
    }
  };
  // copies of the rectangle in s_rect[slot] (view v) into buffer bufi.  Warp = one rectangle row at a time, lane = (texel l >> 3, chunk
  // l & 7): one warp instruction covers 4 whole 128-byte lines of the channels-last copy (an LDGSTS costs one shared-memory wavefront
  // per global line it touches: with one texel per lane it was 27 wavefronts per instruction, profiles/r02p), addresses advance by
  // constants inside a row
  const int st_c = lane & 7, st_t = lane >> 3, wid = t >> 5;
  auto stage = [&](int v, int slot, int bufi) {
    const int4 R = *reinterpret_cast<const int4*>(s_rect[slot]);
    const int rx = R.x, ry = R.y, rw = R.z, rh = R.w;
    const float* src = src_b + ((size_t)(v - 1) * P) * C + 4 * st_c;
    const unsigned dst_c = sbase + bufi * kBufBytes + st_c * kChunkPitch + st_t * 16;
    for (int r = wid; r < rh; r += kThreads / 32) {
      const int y = ry + r;
      const bool row_ok = (unsigned)y < (unsigned)h;
      const float* rowp = src + (size_t)(unsigned)(min(max(y, 0), h - 1) * w * C);
      unsigned dst = dst_c + r * rw * 16;
      for (int tx = st_t; tx < rw; tx += 4, dst += 64) {
        const int x = rx + tx;
        const bool inb = row_ok && ((unsigned)x < (unsigned)w);
        const float* g = rowp + (unsigned)(min(max(x, 0), w - 1) * C);
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(dst), "l"(g), "r"(inb ? 16 : 0) : "memory");   // 0: zero fill
      }
    }
  };

  float2 sum[kCG / 2], sq[kCG / 2];         // sums over the SOURCE views; the key view's terms (:125-126) are added when a group is written
  auto init_acc = [&]() {
#pragma unroll
    for (int k = 0; k < kCG / 2; ++k) sum[k] = sq[k] = make_float2(0.f, 0.f);
  };
  auto gather = [&](float depth, bool plane_ok, int slot, int bufi, int v) {
    float ix, iy;
    homo_coord_fast(s_M[v - 1], xf, yf, depth, h, w, y_w, y_h, ix, iy);
    const float fx = floorf(ix), fy = floorf(iy);
    // ATen zeros padding: a corner outside the map contributes nothing; some corner inside <=> x0 in [-1, w-1], y0 in [-1, h-1]
    const bool needed = plane_ok && (fx >= -1.0f) && (fx <= wm) && (fy >= -1.0f) && (fy <= hm);   // false for nan / inf
    if (!needed) return;
    const int x0 = (int)fx, y0 = (int)fy;
    const float wx = fsub(ix, fx), wy = fsub(iy, fy);
    const float wx0 = fsub(1.0f, wx), wy0 = fsub(1.0f, wy);
    const float nw = fmul(wy0, wx0), ne = fmul(wy0, wx), sw = fmul(wy, wx0), se = fmul(wy, wx);
    const int4 R = *reinterpret_cast<const int4*>(s_rect[slot]);
    const int rw = R.z, rh = R.w;
    const int dx = x0 - R.x, dy = y0 - R.y;
    if ((dx >= 0) && (dx < rw - 1) && (dy >= 0) && (dy < rh - 1)) {   // the whole cell is staged (never when rw = 0)
      const unsigned char* a0 = s_buf + bufi * kBufBytes + (dy * rw + dx) * 16;
      const unsigned char* a1 = a0 + rw * 16;
#pragma unroll
      for (int c = 0; c < kChunks; ++c) {
        const float4 A = *reinterpret_cast<const float4*>(a0 + c * kChunkPitch), Bq = *reinterpret_cast<const float4*>(a0 + c * kChunkPitch + 16);
        const float4 Cq = *reinterpret_cast<const float4*>(a1 + c * kChunkPitch), Dq = *reinterpret_cast<const float4*>(a1 + c * kChunkPitch + 16);
        accumulate4<kPacked>(A, Bq, Cq, Dq, nw, ne, sw, se, sum + 2 * c, sq + 2 * c);
      }
    } else {
      const float* src = src_b + ((size_t)(v - 1) * P) * C;
      const bool xl = x0 >= 0, xr = x0 + 1 < w, yt = y0 >= 0, yb = y0 + 1 < h;
      const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
      const float* g00 = src + ((size_t)max(y0, 0) * w + max(x0, 0)) * C;
      const float* g01 = src + ((size_t)max(y0, 0) * w + min(x0 + 1, w - 1)) * C;
      const float* g10 = src + ((size_t)min(y0 + 1, h - 1) * w + max(x0, 0)) * C;
      const float* g11 = src + ((size_t)min(y0 + 1, h - 1) * w + min(x0 + 1, w - 1)) * C;
#pragma unroll
      for (int c = 0; c < kChunks; ++c) {   // fully unrolled: a dynamic index would move the accumulators to local memory
        const float4 A = (yt && xl) ? ldg4(g00 + 4 * c) : zero, Bq = (yt && xr) ? ldg4(g01 + 4 * c) : zero;
        const float4 Cq = (yb && xl) ? ldg4(g10 + 4 * c) : zero, Dq = (yb && xr) ? ldg4(g11 + 4 * c) : zero;
        accumulate4<kPacked>(A, Bq, Cq, Dq, nw, ne, sw, se, sum + 2 * c, sq + 2 * c);
      }
    }
  };
  const float inv_V = 1.0f / (float)Vt;        // x / V as x * (1 / V): within 1 ulp, far inside the 1e-4 volume tolerance
  const size_t cstride = (size_t)out_D * P;
  auto write_out = [&](int pg) {               // variance = sq / V - (sum / V)^2 (:135), streaming stores
    const int dl = pg * kPL + pl;
    if (!in_img || dl >= Dl) return;
    float* out = variance + (((size_t)b * C + c0) * out_D + out_d0 + dl) * P + p;
    const float2 iv = make_float2(inv_V, inv_V);
#pragma unroll
    for (int k = 0; k < kCG / 2; ++k) {
      const float2 r = s_ref[k * kPx + px];
      const float2 m = __fmul2_rn(__fadd2_rn(sum[k], r), iv);
      const float2 q = __fmul2_rn(__ffma2_rn(r, r, sq[k]), iv);
      __stcs(out, fsub(q.x, fmul(m.x, m.x)));
      __stcs(out + cstride, fsub(q.y, fmul(m.y, m.y)));
      out += 2 * cstride;
    }
  };
  auto advance = [&](int& pg, int& v) {
    if (++v > nV) { v = 1; ++pg; }
  };

  int pg = pg0, v = 1;                         // step s
  int pg1_ = pg0, v1 = 1;                      // step s + 1
  advance(pg1_, v1);
  int pg2 = pg1_, v2 = v1;                     // step s + 2
  advance(pg2, v2);
  if (t < 32) {
    plan(pg, v, 0);
    if (nsteps > 1) plan(pg1_, v1, 1);
  }
  init_acc();
  float depth = __ldg(dep_b + min(pg * kPL + pl, Dl - 1));
  __syncthreads();
  stage(v, 0, 0);
  asm volatile("cp.async.wait_all;\n" ::: "memory");
  __syncthreads();
  int sl0 = 0, sl1 = 1, sl2 = 2;              // s_rect slots of steps s, s + 1, s + 2
  for (int s = 0; s < nsteps; ++s) {
    if (t < 32 && s + 2 < nsteps) plan(pg2, v2, sl2);
    if (s + 1 < nsteps) stage(v1, sl1, (s + 1) & 1);           // in flight while this step is gathered
    const float depth_next = (v == nV) ? __ldg(dep_b + min((pg + 1) * kPL + pl, Dl - 1)) : depth;
    gather(depth, in_img && (pg * kPL + pl < Dl), sl0, s & 1, v);
    if (v == nV) {
      write_out(pg);
      if (s + 1 < nsteps) init_acc();
    }
    depth = depth_next;
    pg = pg1_; v = v1; pg1_ = pg2; v1 = v2;
    { const int o = sl0; sl0 = sl1; sl1 = sl2; sl2 = o; }
    advance(pg2, v2);
    asm volatile("cp.async.wait_all;\n" ::: "memory");
    __syncthreads();                           // next rectangle landed, its successor's plan is visible, this buffer is free
  }
}


// ---------------------------------------------------------------------------------------------------------
// Warp-specialised form of the same kernel: one CTA per SM = 16 consumer warps (tile of 64 key pixels x 8 planes, thread = voxel x 32
// channels) + ONE producer warp that plans and copies the rectangles of the coming steps into a ring of kWsBufs buffers.  The
// consumers execute no planning and no copy instructions and meet no CTA barrier: they wait on the ring's `full` mbarrier (32
// cp.async completions + the producer's release of the rectangle record), gather, and release the buffer through its `empty`
// mbarrier (one arrival per consumer warp).  The two-barriers / one-barrier kernels above lost 20 % of their warp-time at the step
// barrier behind whichever warp had planned and copied (profiles/r02r_var_kernel.md).
// ---------------------------------------------------------------------------------------------------------
constexpr int kWsPL = 8;                           // planes per step
constexpr int kWsConsumers = kPx * kWsPL;          // 512 threads
constexpr int kWsThreads = kWsConsumers + 128;     // + the producer's warpgroup (one working warp; setmaxnreg moves its registers to the consumers)
constexpr int kWsBufs = 4;                         // staged rectangles in flight

template <bool kPacked>
__global__ void __launch_bounds__(kWsThreads, 1) mvs_variance_ws_kernel(
    const float* __restrict__ feats, const float* __restrict__ src_cl, const float* __restrict__ proj, const float* __restrict__ depths,
    int B, int Vt, int C, int h, int w, int D, int d_begin, int d_end, float* __restrict__ variance, int out_D, int out_d0, int tiles_x,
    int ncg, int nsplit) {
  constexpr int kPL = kWsPL, kThreads = kWsThreads;
  static_assert(4 * kPL == 32, "the producer warp plans a step: 4 corner pixels x kPL planes");
  extern __shared__ __align__(128) unsigned char s_buf[];   // kWsBufs staged rectangles
  __shared__ float s_M[kMaxMvsViews][12];
  __shared__ __align__(16) int s_rect[kWsBufs][4];   // {rx, ry, rw, rh} of the step in each buffer; rw = 0: nothing staged
  __shared__ __align__(8) unsigned long long s_bar[2 * kWsBufs];   // full[b]: rectangle landed; empty[b]: every consumer warp is done with it
  __shared__ float2 s_ref[(kCG / 2) * kPx];   // key-view features of the tile, [channel pair][tile pixel]
  const int t = threadIdx.x, lane = t & 31;
  const int b = blockIdx.z;
  const int P = h * w, nV = Vt - 1;
  const int tyi = blockIdx.x / tiles_x, txi = blockIdx.x - tyi * tiles_x;
  const int split = blockIdx.y / ncg, cg = blockIdx.y - split * ncg;
  const int c0 = cg * kCG;
  for (int e = t; e < nV * 12; e += kThreads) {
    const int v = 1 + e / 12, i = (e % 12) / 4, j = e % 4;
    s_M[v - 1][i * 4 + j] = matmul4_elem(proj + ((size_t)b * Vt + v) * 16, proj + ((size_t)b * Vt) * 16, i, j);
  }
  const unsigned bar0 = smem_u32(s_bar);
  if (t == 0) {
    for (int bb = 0; bb < kWsBufs; ++bb) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar0 + 8 * bb), "r"(129));                   // 128 copy completions + 1 release of the plan
      asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar0 + 8 * (kWsBufs + bb)), "r"(kWsConsumers / 32));
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }

  const bool producer = t >= kWsConsumers;
  const int px = t & (kPx - 1), pl = producer ? 0 : t / kPx;   // producer lanes: pl unused
  const int xi = txi * kTW + (px % kTW), yi = tyi * kTH + (px / kTW);
  const bool in_img = (xi < w) && (yi < h);
  const int p = min(yi, h - 1) * w + min(xi, w - 1);
  const int Dl = d_end - d_begin;
  const int npg = (Dl + kPL - 1) / kPL, pg_per = (npg + nsplit - 1) / nsplit;
  const int pg0 = split * pg_per, pg1 = min(npg, pg0 + pg_per);
  if (pg0 >= pg1) return;                                   // uniform
  for (int e = t; e < kCG * kPx; e += kThreads) {           // visible after the first barrier below
    const int k = e / kPx, q = e - k * kPx;
    const int qx = txi * kTW + (q % kTW), qy = tyi * kTH + (q / kTW);
    reinterpret_cast<float*>(s_ref)[((k >> 1) * kPx + q) * 2 + (k & 1)] =
        (qx < w && qy < h) ? __ldg(feats + (((size_t)b * Vt) * C + c0 + k) * P + (size_t)qy * w + qx) : 0.f;
  }
  const int nsteps = (pg1 - pg0) * nV;
  const float y_w = rcp_refined((float)(w - 1) * 0.5f), y_h = rcp_refined((float)(h - 1) * 0.5f);
  const float wm = (float)(w - 1), hm = (float)(h - 1);
  const float xf = (float)xi, yf = (float)yi;
  const unsigned sbase = smem_u32(s_buf);
  const float* src_b = src_cl + ((size_t)b * nV * P) * C + c0;
  const float* dep_b = depths + (size_t)b * D + d_begin;

  // warp 0: rectangle of (plane group pg, view v) -> s_rect[slot]
  auto plan = [&](int pg, int v, int slot) {
    const int corner = lane & 3, cpl = lane >> 2;
    const int dl = pg * kPL + cpl;
    const bool valid = (cpl < kPL) && (dl < Dl);
    const int cx = (corner & 1) ? min(txi * kTW + kTW - 1, w - 1) : txi * kTW;
    const int cy = (corner & 2) ? min(tyi * kTH + kTH - 1, h - 1) : tyi * kTH;
    const float d = __ldg(dep_b + min(dl, Dl - 1));
    const float* M = s_M[v - 1];
    float ix, iy;
    homo_coord_fast(M, (float)cx, (float)cy, d, h, w, y_w, y_h, ix, iy);
    const float den = fadd(__fmaf_rn(M[10], d, __fmaf_rn(M[9], fmul((float)cy, d), fmul(M[8], fmul((float)cx, d)))), M[11]);
    const bool pos = den > 0.f;
    const bool pos0 = __shfl_sync(kFullMask, pos, lane & ~3);
    const bool regular = !valid || (isfinite(ix) && isfinite(iy) && (den != 0.f) && (pos == pos0));
    const bool all_regular = __all_sync(kFullMask, regular);
    // clamp before the conversion; margin 2^-8 px for the rounding differences between a corner's chain and an interior pixel's
    const float lo_x = fminf(fmaxf(floorf(ix - 0.00390625f), -1.f), (float)w), hi_x = fminf(fmaxf(floorf(ix + 0.00390625f) + 1.f, -1.f), (float)w);
    const float lo_y = fminf(fmaxf(floorf(iy - 0.00390625f), -1.f), (float)h), hi_y = fminf(fmaxf(floorf(iy + 0.00390625f) + 1.f, -1.f), (float)h);
    const bool use = valid && all_regular;
    const int xlo = __reduce_min_sync(kFullMask, use ? (int)lo_x : kNone), xhi = __reduce_max_sync(kFullMask, use ? (int)hi_x : -kNone);
    const int ylo = __reduce_min_sync(kFullMask, use ? (int)lo_y : kNone), yhi = __reduce_max_sync(kFullMask, use ? (int)hi_y : -kNone);
    if (lane == 0) {
      int rw = 0, rh = 0;
      // some cell with a corner inside the map: x0 in [-1, w-1] <=> columns [x0, x0 + 1] meet [0, w-1]
      if (all_regular && xlo != kNone && xlo <= w - 1 && xhi >= 0 && ylo <= h - 1 && yhi >= 0) {
        rw = xhi - xlo + 1;
        rh = yhi - ylo + 1;
        if (rw * rh > kCap) rw = rh = 0;
      }
      s_rect[slot][0] = xlo; s_rect[slot][1] = ylo; s_rect[slot][2] = rw; s_rect[slot][3] = rh;
    }
  };
  // copies of the rectangle in s_rect[slot] (view v) into buffer bufi.  Warp = one rectangle row at a time, lane = (texel l >> 3, chunk
  // l & 7): one warp instruction covers 4 whole 128-byte lines of the channels-last copy (an LDGSTS costs one shared-memory wavefront
  // per global line it touches: with one texel per lane it was 27 wavefronts per instruction, profiles/r02p), addresses advance by
  // constants inside a row
  const int st_c = lane & 7, st_t = lane >> 3;
  auto stage = [&](int v, int slot, int bufi, int r0, int rstep) {
    const int4 R = *reinterpret_cast<const int4*>(s_rect[slot]);
    const int rx = R.x, ry = R.y, rw = R.z, rh = R.w;
    const float* src = src_b + ((size_t)(v - 1) * P) * C + 4 * st_c;
    const unsigned dst_c = sbase + bufi * kBufBytes + st_c * kChunkPitch + st_t * 16;
    for (int r = r0; r < rh; r += rstep) {                 // the producer warps share the rows
      const int y = ry + r;
      const bool row_ok = (unsigned)y < (unsigned)h;
      const float* rowp = src + (size_t)(unsigned)(min(max(y, 0), h - 1) * w * C);
      unsigned dst = dst_c + r * rw * 16;
      for (int tx = st_t; tx < rw; tx += 4, dst += 64) {
        const int x = rx + tx;
        const bool inb = row_ok && ((unsigned)x < (unsigned)w);
        const float* g = rowp + (unsigned)(min(max(x, 0), w - 1) * C);
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(dst), "l"(g), "r"(inb ? 16 : 0) : "memory");   // 0: zero fill
      }
    }
  };

  float2 sum[kCG / 2], sq[kCG / 2];         // sums over the SOURCE views; the key view's terms (:125-126) are added when a group is written
  auto init_acc = [&]() {
#pragma unroll
    for (int k = 0; k < kCG / 2; ++k) sum[k] = sq[k] = make_float2(0.f, 0.f);
  };
  auto gather = [&](float depth, bool plane_ok, int slot, int bufi, int v) {
    float ix, iy;
    homo_coord_fast(s_M[v - 1], xf, yf, depth, h, w, y_w, y_h, ix, iy);
    const float fx = floorf(ix), fy = floorf(iy);
    // ATen zeros padding: a corner outside the map contributes nothing; some corner inside <=> x0 in [-1, w-1], y0 in [-1, h-1]
    const bool needed = plane_ok && (fx >= -1.0f) && (fx <= wm) && (fy >= -1.0f) && (fy <= hm);   // false for nan / inf
    if (!needed) return;
    const int x0 = (int)fx, y0 = (int)fy;
    const float wx = fsub(ix, fx), wy = fsub(iy, fy);
    const float wx0 = fsub(1.0f, wx), wy0 = fsub(1.0f, wy);
    const float nw = fmul(wy0, wx0), ne = fmul(wy0, wx), sw = fmul(wy, wx0), se = fmul(wy, wx);
    const int4 R = *reinterpret_cast<const int4*>(s_rect[slot]);
    const int rw = R.z, rh = R.w;
    const int dx = x0 - R.x, dy = y0 - R.y;
    if ((dx >= 0) && (dx < rw - 1) && (dy >= 0) && (dy < rh - 1)) {   // the whole cell is staged (never when rw = 0)
      const unsigned char* a0 = s_buf + bufi * kBufBytes + (dy * rw + dx) * 16;
      const unsigned char* a1 = a0 + rw * 16;
#pragma unroll
      for (int c = 0; c < kChunks; ++c) {
        const float4 A = *reinterpret_cast<const float4*>(a0 + c * kChunkPitch), Bq = *reinterpret_cast<const float4*>(a0 + c * kChunkPitch + 16);
        const float4 Cq = *reinterpret_cast<const float4*>(a1 + c * kChunkPitch), Dq = *reinterpret_cast<const float4*>(a1 + c * kChunkPitch + 16);
        accumulate4<kPacked>(A, Bq, Cq, Dq, nw, ne, sw, se, sum + 2 * c, sq + 2 * c);
      }
    } else {
      const float* src = src_b + ((size_t)(v - 1) * P) * C;
      const bool xl = x0 >= 0, xr = x0 + 1 < w, yt = y0 >= 0, yb = y0 + 1 < h;
      const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
      const float* g00 = src + ((size_t)max(y0, 0) * w + max(x0, 0)) * C;
      const float* g01 = src + ((size_t)max(y0, 0) * w + min(x0 + 1, w - 1)) * C;
      const float* g10 = src + ((size_t)min(y0 + 1, h - 1) * w + max(x0, 0)) * C;
      const float* g11 = src + ((size_t)min(y0 + 1, h - 1) * w + min(x0 + 1, w - 1)) * C;
#pragma unroll
      for (int c = 0; c < kChunks; ++c) {   // fully unrolled: a dynamic index would move the accumulators to local memory
        const float4 A = (yt && xl) ? ldg4(g00 + 4 * c) : zero, Bq = (yt && xr) ? ldg4(g01 + 4 * c) : zero;
        const float4 Cq = (yb && xl) ? ldg4(g10 + 4 * c) : zero, Dq = (yb && xr) ? ldg4(g11 + 4 * c) : zero;
        accumulate4<kPacked>(A, Bq, Cq, Dq, nw, ne, sw, se, sum + 2 * c, sq + 2 * c);
      }
    }
  };
  const float inv_V = 1.0f / (float)Vt;        // x / V as x * (1 / V): within 1 ulp, far inside the 1e-4 volume tolerance
  const size_t cstride = (size_t)out_D * P;
  auto write_out = [&](int pg) {               // variance = sq / V - (sum / V)^2 (:135), streaming stores
    const int dl = pg * kPL + pl;
    if (!in_img || dl >= Dl) return;
    float* out = variance + (((size_t)b * C + c0) * out_D + out_d0 + dl) * P + p;
    const float2 iv = make_float2(inv_V, inv_V);
#pragma unroll
    for (int k = 0; k < kCG / 2; ++k) {
      const float2 r = s_ref[k * kPx + px];
      const float2 m = __fmul2_rn(__fadd2_rn(sum[k], r), iv);
      const float2 q = __fmul2_rn(__ffma2_rn(r, r, sq[k]), iv);
      __stcs(out, fsub(q.x, fmul(m.x, m.x)));
      __stcs(out + cstride, fsub(q.y, fmul(m.y, m.y)));
      out += 2 * cstride;
    }
  };
  auto advance = [&](int& pg, int& v) {
    if (++v > nV) { v = 1; ++pg; }
  };

  auto mbar_wait = [&](unsigned bar, unsigned parity) {
    unsigned done = 0;
    while (!done)
      asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.b32 %0, 1, 0, p;\n\t}\n"
                   : "=r"(done) : "r"(bar), "r"(parity) : "memory");
  };
  __syncthreads();                             // transforms, key tile and barriers are set up; the roles part ways here
  int pg = pg0, v = 1;
  // 640 threads x 96 registers at launch (5 warps per scheduler partition); the producer's warpgroup gives 32 registers per thread
  // back and the four consumer warpgroups take 8 each: 4 x 104 + 64 = 5 x 96
  if (producer) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 64;\n");
    // producer warpgroup (4 warps): per step j, as soon as the consumers have released buffer j % kWsBufs, warp 0 plans the
    // rectangle, the group meets on a named barrier, every warp copies every fourth row
    const int pw = (t - kWsConsumers) >> 5;
    for (int j = 0; j < nsteps; ++j) {
      const int bb = j % kWsBufs, u = j / kWsBufs;
      if (u > 0) mbar_wait(bar0 + 8 * (kWsBufs + bb), (unsigned)(u - 1) & 1u);
      if (pw == 0) plan(pg, v, bb);
      asm volatile("bar.sync 1, 128;\n" ::: "memory");
      stage(v, bb, bb, pw, 4);
      asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(bar0 + 8 * bb) : "memory");   // arrives when this lane's copies have landed
      if (t == kWsConsumers) asm volatile("mbarrier.arrive.release.cta.shared::cta.b64 _, [%0];" ::"r"(bar0 + 8 * bb) : "memory");   // publishes s_rect[bb]
      advance(pg, v);
    }
    asm volatile("cp.async.wait_all;\n" ::: "memory");
    return;
  }
  asm volatile("setmaxnreg.inc.sync.aligned.u32 104;\n");
  init_acc();
  float depth = __ldg(dep_b + min(pg * kPL + pl, Dl - 1));
  for (int s = 0; s < nsteps; ++s) {
    const int bb = s % kWsBufs, u = s / kWsBufs;
    const float depth_next = (v == nV) ? __ldg(dep_b + min((pg + 1) * kPL + pl, Dl - 1)) : depth;
    mbar_wait(bar0 + 8 * bb, (unsigned)u & 1u);
    gather(depth, in_img && (pg * kPL + pl < Dl), bb, bb, v);
    __syncwarp();
    if (lane == 0) asm volatile("mbarrier.arrive.release.cta.shared::cta.b64 _, [%0];" ::"r"(bar0 + 8 * (kWsBufs + bb)) : "memory");
    if (v == nV) {
      write_out(pg);
      if (s + 1 < nsteps) init_acc();
    }
    depth = depth_next;
    advance(pg, v);
  }
}


}  // namespace varcl
}  // namespace rmvd

using namespace rmvd;

extern "C" size_t rmvd_mvs_variance_workspace_bytes(int B, int Vt, int C, int h, int w) {
  if (B <= 0 || Vt <= 1 || C <= 0 || h <= 0 || w <= 0 || C % varcl::kCG != 0) return 0;   // shapes the staged kernel does not take need none
  if ((long long)C * h * w >= (1ll << 31)) return 0;                                       // 32-bit texel offsets inside one map
  return sizeof(float) * (size_t)B * (Vt - 1) * C * h * w;
}

namespace {
template <int kPL, bool kPacked>
int launch_cl(const float* feats, float* ws, const float* proj, const float* depths, int B, int Vt, int C, int h, int w, int D, int d_begin,
              int d_end, float* variance, int out_D, int out_d0, cudaStream_t stream) {
  using namespace rmvd::varcl;
  auto kernel = mvs_variance_cl_kernel<kPL, kPacked>;
  static int sm_count = 0;
  if (sm_count == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev);
    if (sm_count <= 0) sm_count = 148;
  }
  const int smem = 2 * kBufBytes;
  cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (e != cudaSuccess) return (int)e;
  const int P = h * w, nV = Vt - 1;
  dim3 g1((P + 31) / 32, C / 32, B * nV);
  if (g1.z > 65535) return RMVD_E_SHAPE;
  nchw_to_cl_kernel<<<g1, 256, 0, stream>>>(feats, ws, Vt, C, P);
  const int tiles_x = (w + kTW - 1) / kTW, tiles_y = (h + kTH - 1) / kTH, ncg = C / kCG;
  const int npg = (d_end - d_begin + kPL - 1) / kPL;
  // plane splits: enough CTAs for ~12 waves of the resident slots, so that the last wave costs little
  const long long slots = (long long)sm_count * (kPL <= 4 ? 2 : 1);
  const long long base = (long long)tiles_x * tiles_y * ncg * B;
  int nsplit = (int)((12 * slots + base - 1) / base);
  nsplit = nsplit < 1 ? 1 : (nsplit > npg ? npg : nsplit);
  if (const char* s = getenv("RMVD_MVS_VARIANCE_SPLIT")) nsplit = max(1, min(npg, atoi(s)));
  if ((long long)nsplit * ncg > 65535) return RMVD_E_SHAPE;
  dim3 grid((unsigned)(tiles_x * tiles_y), (unsigned)(nsplit * ncg), B);
  kernel<<<grid, kPx * kPL, smem, stream>>>(feats, ws, proj, depths, B, Vt, C, h, w, D, d_begin, d_end, variance, out_D, out_d0, tiles_x, ncg,
                                            nsplit);
  return (int)cudaGetLastError();
}

template <bool kPacked>
int launch_ws(const float* feats, float* ws, const float* proj, const float* depths, int B, int Vt, int C, int h, int w, int D, int d_begin,
              int d_end, float* variance, int out_D, int out_d0, cudaStream_t stream) {
  using namespace rmvd::varcl;
  auto kernel = mvs_variance_ws_kernel<kPacked>;
  static int sm_count = 0;
  if (sm_count == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev);
    if (sm_count <= 0) sm_count = 148;
  }
  const int smem = kWsBufs * kBufBytes;
  cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (e != cudaSuccess) return (int)e;
  const int P = h * w, nV = Vt - 1;
  dim3 g1((P + 31) / 32, C / 32, B * nV);
  if (g1.z > 65535) return RMVD_E_SHAPE;
  nchw_to_cl_kernel<<<g1, 256, 0, stream>>>(feats, ws, Vt, C, P);
  const int tiles_x = (w + kTW - 1) / kTW, tiles_y = (h + kTH - 1) / kTH, ncg = C / kCG;
  const int npg = (d_end - d_begin + kWsPL - 1) / kWsPL;
  const long long base = (long long)tiles_x * tiles_y * ncg * B;
  int nsplit = (int)((6LL * sm_count + base - 1) / base);      // at least ~6 waves of one CTA per SM (measured at cfg3: 1 split 1.008 ms, 2: 1.033, 4: 1.10)
  nsplit = nsplit < 1 ? 1 : (nsplit > npg ? npg : nsplit);
  if (const char* s = getenv("RMVD_MVS_VARIANCE_SPLIT")) nsplit = max(1, min(npg, atoi(s)));
  if ((long long)nsplit * ncg > 65535) return RMVD_E_SHAPE;
  dim3 grid((unsigned)(tiles_x * tiles_y), (unsigned)(nsplit * ncg), B);
  kernel<<<grid, kWsThreads, smem, stream>>>(feats, ws, proj, depths, B, Vt, C, h, w, D, d_begin, d_end, variance, out_D, out_d0, tiles_x, ncg,
                                             nsplit);
  return (int)cudaGetLastError();
}
}  // namespace

extern "C" int rmvd_mvs_variance_fwd(const float* feats, const float* proj, const float* depths, int B, int Vt, int C, int h, int w, int D,
                                     int d_begin, int d_end, float* variance, int out_D, int out_d0, void* stream);

extern "C" int rmvd_mvs_variance_fwd_ws(const float* feats, const float* proj, const float* depths, int B, int Vt, int C, int h, int w, int D,
                                        int d_begin, int d_end, float* variance, int out_D, int out_d0, void* workspace,
                                        size_t workspace_bytes, void* stream) {
  if (!feats || !proj || !depths || !variance) return RMVD_E_NULL;
  if (B <= 0 || Vt <= 0 || C <= 0 || h <= 0 || w <= 0 || D <= 0 || out_D <= 0 || B > 65535) return RMVD_E_SHAPE;
  if (Vt - 1 > kMaxMvsViews) return RMVD_E_SHAPE;
  if (d_begin < 0 || d_end > D || d_begin > d_end || out_d0 < 0 || out_d0 + (d_end - d_begin) > out_D) return RMVD_E_RANGE;
  if (d_begin == d_end) return RMVD_OK;
  const size_t need = rmvd_mvs_variance_workspace_bytes(B, Vt, C, h, w);
  static const int mode = [] { const char* s = getenv("RMVD_MVS_VARIANCE_KERNEL"); return s ? atoi(s) : 0; }();   // A/B switch (development)
  if (need == 0 || mode == 9)   // shapes outside the staged kernel (C not a multiple of 32, a single view): the NCHW-staged kernel of mvs.cu
    return rmvd_mvs_variance_fwd(feats, proj, depths, B, Vt, C, h, w, D, d_begin, d_end, variance, out_D, out_d0, stream);
  if (!workspace) return RMVD_E_NULL;
  if (workspace_bytes < need) return RMVD_E_WORKSPACE;
  if ((uintptr_t)workspace % 16 != 0) return RMVD_E_WORKSPACE;   // 16-byte copies
  float* ws = (float*)workspace;
  cudaStream_t st = (cudaStream_t)stream;
  switch (mode) {   // 0 (default): warp-specialised kernel, packed fp32; the others are the A/B variants of DESIGN.md 5
    case 1: return launch_ws<false>(feats, ws, proj, depths, B, Vt, C, h, w, D, d_begin, d_end, variance, out_D, out_d0, st);
    case 2: return launch_cl<4, true>(feats, ws, proj, depths, B, Vt, C, h, w, D, d_begin, d_end, variance, out_D, out_d0, st);
    case 3: return launch_cl<4, false>(feats, ws, proj, depths, B, Vt, C, h, w, D, d_begin, d_end, variance, out_D, out_d0, st);
    case 4: return launch_cl<8, true>(feats, ws, proj, depths, B, Vt, C, h, w, D, d_begin, d_end, variance, out_D, out_d0, st);
    default: return launch_ws<true>(feats, ws, proj, depths, B, Vt, C, h, w, D, d_begin, d_end, variance, out_D, out_d0, st);
  }
}
