// Plane-sweep correlation, persistent warp-specialised form (sm_100a): one CTA per SM owns all 512 TMEM columns and
// loops over work items (tile of 64 key pixels, sample, source view).  The contraction:
// corr = bilinear(G), G = key . band^T over the source texels under the tile's epipolar band
// (planesweep_corr.py:174-186 restricted to where it is sampled), bf16 hi/lo split operands on tcgen05.  The serial
// phases of a tile do not wait for one another:
//
//   warp 0      planner    per item: calibration (planesweep_corr.py:228-300), tile slots; per chunk of planes: the band
//                          (one 8-aligned [xa, xa + len) span per source row), the per-plane terms and dead-plane bits,
//                          published as a self-contained chunk descriptor in a ring of three
//   warp 1      producer   per (pass of 128 band texels, 32 channels): one bulk copy (cp.async.bulk, no tensor map) per
//                          band row straight into the UMMA no-swizzle K-major stage; ring of six 16 KB stages
//   warp 2      MMA        one elected lane: 4 x tcgen05.mma (M = 128 = [key_hi ; key_lo] rows, N <= 128, K = 16) per stage
//                          into one of three 128-column accumulators; tcgen05.commit frees stages / publishes passes
//   warps 4-15  consumers  drain the accumulators (tcgen05.ld, key_hi + key_lo rows added by one shuffle) into the shared
//                          G tile, then sample it: filtered locate (integer corner and 0/1 mask bit-exact, see below),
//                          four G reads, 128-byte streaming stores of corr and mask.  While they sample chunk i the
//                          planner / producer / MMA warps already work on chunk i + 1 (or the next tile); the key tile
//                          of the next item goes to TMEM during the last chunk of the current one.
//
// Key operand layout ("tile-ordered split", written by nchw_to_keytiles_kernel): per (sample, tile) C / 16 blocks of 4 KB, block ks =
// the 128 A' rows (TMEM lane 32 q + l = key_hi of tile slot 16 q + l for l < 16, key_lo of slot 16 q + l - 16 for l >= 16) x 16
// channels as [8-row group 0..15][chunk 0..1][row 0..7][8 bf16]: the canonical layout tcgen05.cp.128x256b copies into 8 TMEM
// columns (tools/utccp_probe.cu).  A tile's key therefore travels through the same stage ring as the band (one bulk copy per
// 16 KB) and reaches TMEM without touching a register; the copy is issued by the MMA thread in order with its MMAs.
//
// Source operand layout ("grouped split", written by nchw_to_grouped_kernel): per (sample, 32-channel block kc, group of 8
// consecutive texels of one source row) one 1024-byte block [chunk j = 0..7][texel r = 0..7][8 bf16]; chunks 0..3 hold
// the hi halves of channels 8j..8j+7, chunks 4..7 the lo halves (x = hi + lo to 2^-17).  Rows are padded to a multiple of
// 8 texels (zeros).  A band row span that starts at a multiple of 8 is therefore ONE contiguous run, and the copy lands as
// UMMA's canonical K-major SWIZZLE_NONE layout (core matrix = 8 texels x 16 bytes; LBO = 128 B between K chunks,
// SBO = 1024 B between 8-texel groups).
#include <cuda.h>
#include <cuda_bf16.h>

#include <cstdio>
#include <cstdlib>

#include "psweep.cuh"

namespace rmvd {
namespace ws {

constexpr int kThreads = 640;                // 4 control warps (planner, producer, MMA, spare) + 16 consumer warps
constexpr int kPx = 64;                      // key pixels per tile (A' has 2 * kPx = 128 rows: key_hi, key_lo)
constexpr int kNT = 128;                     // band texels per pass = N of one MMA = columns of one accumulator
constexpr int kNBuf = 3;                     // accumulators: a whole chunk (<= 3 passes) can sit in TMEM
constexpr int kKC = 32;                      // channels per stage
#ifndef WS_STAGES
#define WS_STAGES 4
#endif
#ifndef WS_RMAX
#define WS_RMAX 512
#endif
constexpr int kStages = WS_STAGES;           // ring of B' stages
constexpr int kStageBytes = kNT * 128;       // 16 KB: 128 texels x [hi 32 ch | lo 32 ch] bf16
constexpr int kMaxC = 256;                   // A' in TMEM: C / 2 columns
constexpr int kTmemA = kNBuf * kNT;          // TMEM columns: D = [0, 384), A' = [384, 384 + C / 2)
constexpr int kTmemCols = 512;
constexpr int kRMax = WS_RMAX;               // band capacity of one chunk (G columns), <= kNBuf * kNT
// G row pitch = 2 (mod 32).  Sampling: lane = pixel = G row, and adjacent pixels read adjacent columns, so a warp's addresses
// step by pitch + 1 = 3 (mod 32): 32 distinct banks (an odd pitch gives a step of 2 and two wavefronts per read).  Drain:
// sixteen rows at one column step by 2 -- the lower half-warp covers the even banks, the upper half (next column) the odd.
constexpr int kGPitch = kRMax + 2;
static_assert(kGPitch % 32 == 2, "G pitch must be 2 mod 32");
constexpr int kMaxRows = 64;                 // source rows a band may span
constexpr int kMaxPlanes = 256;              // planes per chunk
constexpr int kDescs = 3;                    // chunk descriptors in flight
constexpr int kSchedMax = 1024;              // planes of the inverse-depth schedule the planner keeps in shared memory
constexpr int kConsWarps = 16;
constexpr int kRegsControl = 64, kRegsConsumer = 104;   // setmaxnreg moves registers inside the CTA's own allocation (20 warps x 96): 4 * 64 + 16 * 104 = 1920
constexpr int kConsThreads = kConsWarps * 32;
constexpr int kStep = kConsWarps / 2;        // planes between two sampling steps of a warp

constexpr int kMaxPeers = RMVD_MAX_PEERS;
struct PeerArgs {                  // depth-slab mode (n > 0): peer-mapped output tensors and the end-of-kernel handshake
  int n, rank;
  float* corr[kMaxPeers];          // [V,N,out_S,h,w] of every rank (symmetric memory), own rank included
  float* mask[kMaxPeers];
  unsigned* flags[kMaxPeers];      // every rank's flag array [n] (symmetric memory): slot r = last epoch rank r completed
  unsigned epoch;                  // this call's epoch (monotone per communicator)
  unsigned* done_ctas;             // local counter of finished CTAs
  float* corr_mc; float* mask_mc;  // multicast addresses of the output tensors, or NULL
  unsigned* q_head;                // push mode: queue header {total, final epoch, pusher CTAs done, -}, then cap flags, then cap records
  unsigned q_cap;
  int item_stride, item_offset;    // this launch works on items j * item_stride + item_offset (tile partition between the ranks); 1, 0 otherwise
};

struct __align__(16) PlaneTerm { float ed, id, md, zs; };   // e*d, i*d, m*d, 1/d of one plane (planesweep_corr.py:334,343,492)
// Texels [xa, xa + len) of one source row of the band; xa and len are multiples of 8.  G column of texel x = offp + x.
struct __align__(16) RowSpan { int xa, lenm1, offp, len; };

enum { kStop = 0, kGemm = 1, kDirect = 2 };
enum { kModePlain = 0, kModePeers = 1, kModeDebug = 2, kModePush = 3 };   // kernel variants: outputs only / + peer stores (depth slabs) / + debug outputs /
                                                                          // local stores + a record of every finished chunk for the pusher kernel

struct __align__(16) ChunkDesc {
  RowSpan rows[kMaxRows];
  PlaneTerm planes[kMaxPlanes];
  int pix[kPx];                    // key pixel (y * w + x) of every tile slot, -1 for empty slots
  unsigned dead[kMaxPlanes / 32];  // planes without any valid sample in this tile
  Epi E;
  const __nv_bfloat16* key;        // tile-ordered split key block of this (sample, tile)
  const __nv_bfloat16* src;        // grouped split source map of the (view, sample)
  unsigned long long out_base;     // element offset of plane index 0 of this (view, sample) in the outputs
  int hs, ws, wg, G;               // wg: 8-texel groups per source row, G: groups per map
  int n, v, kind, first, last;     // first / last chunk of its tile
  int s0, len, R, ymin, H, out_plane0;
  int pad_[1];
};
static_assert(sizeof(ChunkDesc) % 16 == 0, "descriptor ring alignment");

struct WsConst {                   // per-launch constants of the out-of-line paths
  float* corr; float* mask; int32_t* dbg_ix0; int32_t* dbg_iy0; uint8_t* dbg_geom;
  const float* G;
  int P, C, w;
  float inv_scale;
  int npeers;                      // depth-slab mode: every output row also goes to the peers' tensors (NVLink stores)
  float* pcorr[kMaxPeers]; float* pmask[kMaxPeers];
  float* mc_corr; float* mc_mask;  // multicast addresses (NVLS): one multimem.st lands in every rank's tensor
};
__shared__ WsConst s_k;

// One output sample.  Single GPU: streaming stores.  Depth slabs over several GPUs (partition 1): the same value is
// stored into the [V,N,S,h,w] tensor of every rank (own rank included) through the peer-mapped pointers, i.e. the all-gather of the
// slabs happens row by row from inside the sampling loop instead of as a separate collective.
// streaming store of base[off]: one 32 x 32 + 64-bit multiply-add forms the address (the compiler's own 64-bit offset arithmetic
// costs four instructions per store in the sampling loop)
__device__ __forceinline__ void stcs_at(float* base, unsigned off, float v) {
  unsigned long long a;
  asm("mad.wide.u32 %0, %1, 4, %2;" : "=l"(a) : "r"(off), "l"(base));
  asm volatile("st.global.cs.f32 [%0], %1;" ::"l"(a), "f"(v) : "memory");
}
__device__ __forceinline__ void stcs_at(float* base, size_t off, float v) { __stcs(base + off, v); }

// corr / mask: the caller's register copies of W.corr / W.mask, already advanced by `base` elements; o: offset from there
template <int kMode, typename Off>
__device__ __forceinline__ void store_out(const WsConst& W, float* __restrict__ corr, float* __restrict__ mask, size_t base, Off o, float c,
                                          float m) {
  if (kMode == kModePush) {      // the pusher kernel re-reads these rows within microseconds: keep them in L2 (no evict-first hint)
    corr[o] = c;
    mask[o] = m;
  } else if (kMode != kModePeers) {
    stcs_at(corr, o, c);
    stcs_at(mask, o, m);
  } else {
    if (W.mc_corr) {             // NVSwitch multicast: the switch replicates the store, this SM sends each value once
      asm volatile("multimem.st.relaxed.sys.global.f32 [%0], %1;\n" ::"l"(W.mc_corr + base + o), "f"(c) : "memory");
      asm volatile("multimem.st.relaxed.sys.global.f32 [%0], %1;\n" ::"l"(W.mc_mask + base + o), "f"(m) : "memory");
    } else {
      for (int g = 0; g < W.npeers; ++g) { W.pcorr[g][base + o] = c; W.pmask[g][base + o] = m; }
    }
  }
}

// barrier slots
constexpr int kBarFull = 0, kBarFree = kBarFull + kStages, kBarAcc = kBarFree + kStages, kBarDrained = kBarAcc + kNBuf,
              kBarCReady = kBarDrained + kNBuf, kBarCFree = kBarCReady + kDescs, kBarAFree = kBarCFree + kDescs,
              kBarJobReady = kBarAFree + 1, kBarJobDone = kBarJobReady + 1, kNumBars = kBarJobDone + 1;

__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
#ifdef RMVD_WS_WATCHDOG
// Development aid: a wait that does not complete within ~2^22 polls reports itself and releases every waiter of the grid, so a
// protocol error ends as a wrong result with a message instead of a hung GPU.
__device__ int g_ws_abort;
#endif
__device__ __forceinline__ void mbar_wait(unsigned bar, unsigned parity) {
  unsigned done = 0;
#ifdef RMVD_WS_WATCHDOG
  unsigned polls = 0;
#endif
  while (!done) {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.b32 %0, 1, 0, p;\n\t}\n"
                 : "=r"(done) : "r"(bar), "r"(parity) : "memory");
#ifdef RMVD_WS_WATCHDOG
    if (!done && (++polls & 1023u) == 0) {
      if (*reinterpret_cast<volatile int*>(&g_ws_abort)) return;
      if (polls > (1u << 22)) {
        if ((threadIdx.x & 31) == 0 || threadIdx.x == 64)
          printf("[psweep_ws watchdog] block %d warp %d: barrier at smem %u parity %u never completed\n", blockIdx.x, threadIdx.x >> 5, bar, parity);
        atomicExch(&g_ws_abort, 1);
        return;
      }
    }
#endif
  }
}
#ifdef RMVD_WS_STATS
// Development build: per-role cycle counters (one recording lane per role), printed by launch_psweep_ws after a synchronise.
__device__ unsigned long long g_ws_stats[32];
#define WS_T0() const long long ws_t0_ = clock64()
#define WS_ADD(i) atomicAdd(&g_ws_stats[i], (unsigned long long)(clock64() - ws_t0_))
#define WS_CNT(i, n) atomicAdd(&g_ws_stats[i], (unsigned long long)(n))
#else
#define WS_T0()
#define WS_ADD(i)
#define WS_CNT(i, n)
#endif
__device__ __forceinline__ void mbar_arrive(unsigned bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];\n" ::"r"(bar) : "memory");
}
__device__ __forceinline__ bool elect_one() {
  unsigned pred;
  asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.u32 %0, 1, 0, P;\n\t}\n" : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void cons_bar() { asm volatile("bar.sync 1, %0;\n" ::"n"(kConsThreads) : "memory"); }
// K-major SWIZZLE_NONE operand: lbo = bytes between 16-byte K chunks, sbo = bytes between 8-row groups; version 1.
__device__ __forceinline__ unsigned long long umma_desc(unsigned addr, unsigned lbo, unsigned sbo) {
  return (unsigned long long)((addr >> 4) & 0x3fffu) | ((unsigned long long)((lbo >> 4) & 0x3fffu) << 16) |
         ((unsigned long long)((sbo >> 4) & 0x3fffu) << 32) | (1ull << 46);
}
__device__ __forceinline__ unsigned umma_idesc_bf16(int M, int N) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((unsigned)(N >> 3) << 17) | ((unsigned)(M >> 4) << 24);
}
__device__ __forceinline__ void tc_mma(unsigned d_tmem, unsigned a_tmem, unsigned long long b_desc, unsigned idesc, unsigned acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
               "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}\n" ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(acc)
               : "memory");
}
__device__ __forceinline__ void tmem_ld32(unsigned addr, float* v) {
  unsigned r[32];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,"
      "%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
        "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),
        "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
        "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(addr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}
// One contiguous run global -> shared, completing its bytes on the mbarrier (SASS: UBLKCP).
__device__ __forceinline__ void bulk_g2s(unsigned dst, const void* src, unsigned bytes, unsigned bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n" ::"r"(dst), "l"(src),
               "r"(bytes), "r"(bar)
               : "memory");
}

// key tile, tile-ordered split layout: element (tile slot, channel c) of the hi rows in bf16 units; the lo row of the same slot sits
// 16 rows (two 8-row groups = 256 elements) further
__device__ __forceinline__ int key_index(int slot, int c) {
  const int r = 32 * (slot >> 4) + (slot & 15);
  return (c >> 4) * 2048 + (r >> 3) * 128 + ((c >> 3) & 1) * 64 + (r & 7) * 8 + (c & 7);
}
__device__ __forceinline__ float key_at(const __nv_bfloat16* key, int slot, int c) {
  const int o = key_index(slot, c);
  return __bfloat162float(key[o]) + __bfloat162float(key[o + 256]);
}
// source map, grouped split layout: element (y, x, c) in bf16 units; the lo half sits 256 elements (4 chunks) further
__device__ __forceinline__ size_t src_index(int G, int wg, int y, int x, int c) {
  return ((size_t)(c >> 5) * G + (size_t)y * wg + (x >> 3)) * 512 + ((c >> 3) & 3) * 64 + (x & 7) * 8 + (c & 7);
}
__device__ __forceinline__ float src_at(const ChunkDesc* D, int y, int x, int c) {
  const size_t o = src_index(D->G, D->wg, y, x, c);
  return __bfloat162float(D->src[o]) + __bfloat162float(D->src[o + 256]);
}

// ---------------------------------------------------------------------------------------------------------
// NCHW fp32 -> grouped split copy of up to kMaxViews source maps in one launch.  CTA = 32 texels of one source row x 128
// channels (grid: row tiles x channel blocks x (map, sample)); kNorm: L2-normalised along C first (normalize(x, dim=1),
// planesweep_corr.py:8-10, :170-172), the CTA then walks all channel blocks itself.
// ---------------------------------------------------------------------------------------------------------
constexpr int kRpCh = 128;
struct GroupBatch {
  const float* src[kMaxViews];
  __nv_bfloat16* dst[kMaxViews];
  int hs[kMaxViews], ws[kMaxViews];
  unsigned* counter;               // work-item counter of the plane-sweep launch that follows: zeroed here
};

// kAct: the maps hold PRE-activation values; LeakyReLU(slope) (x > 0 ? x : x * slope, as ATen computes it) is applied as they are read
template <bool kNorm, bool kAct>
__global__ void __launch_bounds__(256) nchw_to_grouped_kernel(const GroupBatch B, int C, int N, float slope) {
  __shared__ float tile[kRpCh][33];
  __shared__ float s_norm[32];
  const int which = blockIdx.z / N, n = blockIdx.z - which * N;
  if (blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && threadIdx.x == 0) { B.counter[0] = 0u; B.counter[1] = 0u; B.counter[2] = 0u; }
  const int hs = B.hs[which], ws = B.ws[which];
  const int wg = (ws + 7) >> 3, tpr = (wg + 3) >> 2;        // groups per row, 32-texel tiles per row
  if ((int)blockIdx.x >= tpr * hs) return;                   // maps of different sizes share the grid of the largest
  const int y = blockIdx.x / tpr, x0 = (blockIdx.x - y * tpr) * 32;
  const int G = hs * wg;
  const size_t Ps = (size_t)hs * ws;
  const float* __restrict__ src = B.src[which] + (size_t)n * C * Ps + (size_t)y * ws;
  __nv_bfloat16* __restrict__ dst = B.dst[which] + (size_t)n * (C >> 5) * G * 512;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const bool in = x0 + tx < ws;
  float inv_norm = 1.0f;
  if (kNorm) {
    float acc = 0.f;
    for (int c = ty; c < C; c += 8) {
      float v = in ? __ldg(src + (size_t)c * Ps + x0 + tx) : 0.f;
      if (kAct) v = v > 0.f ? v : __fmul_rn(v, slope);
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
    inv_norm = s_norm[tx];                                    // divisor of this thread's loads (pixel tx)
  }
  const int cb0 = kNorm ? 0 : blockIdx.y * kRpCh, cb1 = kNorm ? C : min(C, cb0 + kRpCh);
  for (int c0 = cb0; c0 < cb1; c0 += kRpCh) {
    const int nc = min(kRpCh, C - c0);                        // multiple of 32
    float v[kRpCh / 8];
#pragma unroll
    for (int k = 0; k < kRpCh / 8; ++k) v[k] = (ty + 8 * k < nc && in) ? __ldg(src + (size_t)(c0 + ty + 8 * k) * Ps + x0 + tx) : 0.f;
    if (kAct) {
#pragma unroll
      for (int k = 0; k < kRpCh / 8; ++k) v[k] = v[k] > 0.f ? v[k] : __fmul_rn(v[k], slope);
    }
#pragma unroll
    for (int k = 0; k < kRpCh / 8; ++k) tile[ty + 8 * k][tx] = kNorm ? fdiv(v[k], inv_norm) : v[k];
    __syncthreads();
    // one 16-byte chunk pair (hi, lo) per item: (channel block kcl, group g, chunk jj, texel r); a warp = the four hi chunks
    // (512 contiguous bytes) and the four lo chunks of one group
    for (int idx = threadIdx.x; idx < (nc >> 5) * 128; idx += 256) {
      const int r = idx & 7, jj = (idx >> 3) & 3, g = (idx >> 5) & 3, kcl = idx >> 7;
      const int gx = (x0 >> 3) + g;
      if (gx < wg) {
        float f[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) f[e] = tile[kcl * 32 + jj * 8 + e][g * 8 + r];
        uint4 hi4, lo4;
        unsigned* hp = reinterpret_cast<unsigned*>(&hi4);
        unsigned* lp = reinterpret_cast<unsigned*>(&lo4);
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const __nv_bfloat162 hi = __floats2bfloat162_rn(f[2 * e], f[2 * e + 1]);
          const __nv_bfloat162 lo = __floats2bfloat162_rn(f[2 * e] - __low2float(hi), f[2 * e + 1] - __high2float(hi));
          hp[e] = *reinterpret_cast<const unsigned*>(&hi);
          lp[e] = *reinterpret_cast<const unsigned*>(&lo);
        }
        __nv_bfloat16* o = dst + ((size_t)((c0 >> 5) + kcl) * G + (size_t)y * wg + gx) * 512 + jj * 64 + r * 8;
        *reinterpret_cast<uint4*>(o) = hi4;
        *reinterpret_cast<uint4*>(o + 256) = lo4;
      }
    }
    __syncthreads();
  }
}

// ---------------------------------------------------------------------------------------------------------
// NCHW fp32 key map -> tile-ordered split copy.  CTA = one tile x 64 channels (one 16 KB stage of the plane-sweep kernel);
// kNorm: L2-normalised along C first, the CTA then walks all channel blocks itself.
// ---------------------------------------------------------------------------------------------------------
// kAct: `key` holds PRE-activation values; LeakyReLU(slope) is applied as they are read and, if act_out is given, the activated map is
// also written there in NCHW float32 (the key map has other consumers: context encoder, decoder skip connection).
template <bool kNorm, bool kAct>
__global__ void __launch_bounds__(256) nchw_to_keytiles_kernel(const float* __restrict__ key, __nv_bfloat16* __restrict__ dst, int C, int h, int w,
                                                               int tw, int th, int tiles_x, float slope, float* __restrict__ act_out) {
  __shared__ float tile[64][66];                    // [channel][tile slot]; pitch 66: conflict-free chunk reads below
  __shared__ float s_norm[64];
  const int tiles = gridDim.x, tl = blockIdx.x, n = blockIdx.z;
  const int by = tl / tiles_x, bx = tl - by * tiles_x;
  const size_t P = (size_t)h * w;
  const float* __restrict__ src = key + (size_t)n * C * P;
  __nv_bfloat16* __restrict__ out = dst + ((size_t)n * tiles + tl) * (size_t)C * 128;   // C * 256 bytes per tile
  const int t = threadIdx.x;
  // this thread's tile slot for the loads: slot = t & 63, channels (t >> 6) + 4 k
  const int slot = t & 63, sy = slot / tw, sx = slot - sy * tw;
  const int py = by * th + sy, pxx = bx * tw + sx;
  const bool in = (sy < th) && (py < h) && (pxx < w);
  const size_t poff = in ? (size_t)py * w + pxx : 0;
  float nrm = 1.0f;
  if (kNorm) {
    float acc = 0.f;
    for (int c = t >> 6; c < C; c += 4) {
      float v = in ? __ldg(src + (size_t)c * P + poff) : 0.f;
      if (kAct) v = v > 0.f ? v : __fmul_rn(v, slope);
      acc = fmaf(v, v, acc);
    }
    tile[t >> 6][slot] = acc;
    __syncthreads();
    if (t < 64) s_norm[t] = sqrtf(tile[0][t] + tile[1][t] + tile[2][t] + tile[3][t]) + 1e-9f;
    __syncthreads();
    nrm = s_norm[slot];
  }
  const int cb0 = kNorm ? 0 : blockIdx.y * 64, cb1 = kNorm ? C : min(C, cb0 + 64);
  for (int c0 = cb0; c0 < cb1; c0 += 64) {
    const int nc = min(64, C - c0);                 // multiple of 32
    float v[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) {
      const int c = (t >> 6) + 4 * k;
      v[k] = (c < nc && in) ? __ldg(src + (size_t)(c0 + c) * P + poff) : 0.f;
    }
    if (kAct) {
#pragma unroll
      for (int k = 0; k < 16; ++k) {
        const int c = (t >> 6) + 4 * k;
        v[k] = v[k] > 0.f ? v[k] : __fmul_rn(v[k], slope);
        if (act_out && c < nc && in) act_out[(size_t)n * C * P + (size_t)(c0 + c) * P + poff] = v[k];
      }
    }
#pragma unroll
    for (int k = 0; k < 16; ++k) tile[(t >> 6) + 4 * k][slot] = kNorm ? fdiv(v[k], nrm) : v[k];
    __syncthreads();
    // 16-byte chunks of the block in output order: idx = ((ks * 16 + g) * 2 + chunk) * 8 + rr
    for (int idx = t; idx < (nc >> 4) * 256; idx += 256) {
      const int rr = idx & 7, chunk = (idx >> 3) & 1, g = (idx >> 4) & 15, ks = idx >> 8;
      const int r = 8 * g + rr, l = r & 31;
      const int sl = 16 * (r >> 5) + (l & 15);
      const bool lo_row = l >= 16;
      uint4 o4;
      unsigned* op = reinterpret_cast<unsigned*>(&o4);
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float a = tile[ks * 16 + chunk * 8 + 2 * e][sl], b = tile[ks * 16 + chunk * 8 + 2 * e + 1][sl];
        const __nv_bfloat162 hi = __floats2bfloat162_rn(a, b);
        const __nv_bfloat162 lo = __floats2bfloat162_rn(a - __low2float(hi), b - __high2float(hi));
        op[e] = lo_row ? *reinterpret_cast<const unsigned*>(&lo) : *reinterpret_cast<const unsigned*>(&hi);
      }
      *reinterpret_cast<uint4*>(out + (size_t)(c0 >> 4) * 2048 + (size_t)idx * 8) = o4;
    }
    __syncthreads();
  }
}

// ---------------------------------------------------------------------------------------------------------
// Exact sample location (bit-identical to epi_sample() + pix_to_unnorm() + bilinear_setup() of common.cuh).
// ---------------------------------------------------------------------------------------------------------
struct Sample {
  int x0, y0;             // floor of the un-normalised coordinate, saturated like sat_index()
  float nw, ne, sw, se;   // bilinear weights, 0 for corners outside the source map
  bool valid, geom;
};
struct Pixel {            // per-thread constants of its key pixel
  float u_h, v_h, k_h, z_pole;
  int mode;               // 1: k_h > 0, 2: k_h < 0, 3: k_h == 0 && m > 0, 0: never visible (planesweep_corr.py:500-504)
};

__device__ __forceinline__ void finish(float ix, float iy, int hs, int ws, bool geom, Sample& L) {
  const float x0f = floorf(ix), y0f = floorf(iy);
  const float wx1 = fsub(ix, x0f), wx0 = fsub(1.0f, wx1);
  const float wy1 = fsub(iy, y0f), wy0 = fsub(1.0f, wy1);
  L.x0 = (int)fminf(fmaxf(x0f, -2.0f), (float)ws);          // saturating keeps every range test exact; NaN -> -2
  L.y0 = (int)fminf(fmaxf(y0f, -2.0f), (float)hs);
  const bool x0_in = (unsigned)L.x0 < (unsigned)ws, x1_in = (unsigned)(L.x0 + 1) < (unsigned)ws;
  const bool y0_in = (unsigned)L.y0 < (unsigned)hs, y1_in = (unsigned)(L.y0 + 1) < (unsigned)hs;
  L.nw = (y0_in && x0_in) ? fmul(wy0, wx0) : 0.0f;
  L.ne = (y0_in && x1_in) ? fmul(wy0, wx1) : 0.0f;
  L.sw = (y1_in && x0_in) ? fmul(wy1, wx0) : 0.0f;
  L.se = (y1_in && x1_in) ? fmul(wy1, wx1) : 0.0f;
  const float msum = fadd(fadd(fadd(L.nw, L.ne), L.sw), L.se);
  L.geom = geom;
  L.valid = sampling_mask(msum) && geom;
}

__device__ __forceinline__ void locate(const Pixel& px, const PlaneTerm& pt, int hs, int ws, Sample& L) {
  const float den = fadd(px.k_h, pt.md), nu = fadd(px.u_h, pt.ed), nv = fadd(px.v_h, pt.id);  // :334, :343
  const float zs = pt.zs;
  const bool vis_src = (px.mode == 1) ? (zs > px.z_pole) : (px.mode == 2) ? (zs < px.z_pole) : (px.mode == 3);
  const bool geom = (zs > 0.0f) && vis_src;                                                     // :499-506
  const float us = replace_nonfinite(fdiv(nu, den)), vs = replace_nonfinite(fdiv(nv, den));
  finish(pix_to_unnorm(us, ws), pix_to_unnorm(vs, hs), hs, ws, geom, L);
}

__device__ __forceinline__ Pixel pixel_consts(const Epi& E, int p, int w) {
  const int yi = p / w, xi = p - yi * w;
  const EpiPix ep = epi_pixel(E, xi, yi);
  Pixel px;
  px.u_h = ep.u_h; px.v_h = ep.v_h; px.k_h = ep.k_h;
  px.z_pole = -fdiv(E.m, ep.k_h);                                                               // :330
  px.mode = (ep.k_h > 0.0f) ? 1 : (ep.k_h < 0.0f) ? 2 : ((ep.k_h == 0.0f) && (E.m > 0.0f)) ? 3 : 0;
  return px;
}

// ---------------------------------------------------------------------------------------------------------
// Band of one (tile, inverse-depth interval [dmin, dmax]): union over the tile's pixels of their dilated epipolar segments,
// one span per source row, start and length multiples of 8 texels (= whole 1024-byte groups of the source layout).
// Executed by ONE converged warp (lane l owns tile slots l, l + 32 and band rows l, l + 32).  Writes rows[0..H); returns R
// (G columns), -1 if a pixel is irregular (pole inside the interval, non-finite end points), or a value > kRMax.
// ---------------------------------------------------------------------------------------------------------
__device__ int band_warp(RowSpan* rows, const int* pix, const Epi& E, int w_key, float dmin, float dmax, int hs, int ws, int& ymin_out,
                         int& H_out) {
  const int lane = threadIdx.x & 31;
  constexpr float kEps = 0.05f;
  float ua[2], va[2], ub[2], vb[2];
  bool act[2], bad = false;
  ymin_out = 0; H_out = 0;
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    const int p = pix[lane + 32 * j];
    act[j] = p >= 0;
    ua[j] = va[j] = ub[j] = vb[j] = 0.f;
    if (act[j]) {
      const int yi = p / w_key, xi = p - yi * w_key;
      const EpiPix e = epi_pixel(E, xi, yi);
      const float dena = e.k_h + E.m * dmin, denb = e.k_h + E.m * dmax;
      bool ok = (dena > 0.f && denb > 0.f) || (dena < 0.f && denb < 0.f);   // no pole inside the interval (false for nan too)
      // approximate reciprocals (relative error 2^-22): far inside the 0.05-texel dilation for coordinates near the map (< 32768),
      // irrelevant beyond it (the spans are clipped to the map)
      const float ra = __fdividef(1.0f, dena), rb = __fdividef(1.0f, denb);
      ua[j] = (e.u_h + E.e * dmin) * ra; va[j] = (e.v_h + E.i * dmin) * ra;
      ub[j] = (e.u_h + E.e * dmax) * rb; vb[j] = (e.v_h + E.i * dmax) * rb;
      ok = ok && (fabsf(ua[j]) < 1e7f) && (fabsf(va[j]) < 1e7f) && (fabsf(ub[j]) < 1e7f) && (fabsf(vb[j]) < 1e7f);
      bad = bad || !ok;
    }
  }
  if (__any_sync(kFullMask, bad)) return -1;
  int ylo[2], yhi[2];
  int wy0 = 0x7fffffff, wy1 = -0x7fffffff;
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    ylo[j] = 1; yhi[j] = 0;
    if (act[j]) {
      ylo[j] = max(0, (int)floorf(fminf(va[j], vb[j]) - 0.5f - kEps));
      yhi[j] = min(hs - 1, (int)floorf(fmaxf(va[j], vb[j]) - 0.5f + kEps) + 1);
      if (ylo[j] <= yhi[j]) { wy0 = min(wy0, ylo[j]); wy1 = max(wy1, yhi[j]); }
    }
  }
  wy0 = __reduce_min_sync(kFullMask, wy0);
  wy1 = __reduce_max_sync(kFullMask, wy1);
  if (wy0 > wy1) return 0;                          // no sample of the tile touches the source map
  const int H = wy1 - wy0 + 1;
  if (H > kMaxRows) return kRMax + H;
  int sxa[2] = {0, 0}, slen[2] = {0, 0};          // lane r keeps band rows r and r + 32
  float inv_dv[2], du[2];
  bool sloped[2];
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    const float dv = vb[j] - va[j];
    du[j] = ub[j] - ua[j];
    sloped[j] = fabsf(dv) > 1e-6f;
    inv_dv[j] = sloped[j] ? __fdividef(1.0f, dv) : 0.f;
  }
  int total = 0;                                    // columns so far (warp-uniform): a band that cannot fit stops early
  for (int y = wy0; y <= wy1; ++y) {               // warp-uniform bounds
    int xlo = 0x7fffffff, xhi = -0x7fffffff;
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      if (act[j] && y >= ylo[j] && y <= yhi[j]) {
        float t0 = 0.f, t1 = 1.f;
        if (sloped[j]) {
          const float ta = ((float)y - 0.5f - kEps - va[j]) * inv_dv[j], tb = ((float)y + 1.5f + kEps - va[j]) * inv_dv[j];
          t0 = fmaxf(0.f, fminf(ta, tb) - 1e-4f);
          t1 = fminf(1.f, fmaxf(ta, tb) + 1e-4f);
        }
        if (t0 <= t1) {
          const float u0 = ua[j] + t0 * du[j], u1 = ua[j] + t1 * du[j];
          const int a = max(0, (int)floorf(fminf(u0, u1) - 0.5f - kEps));
          const int b = min(ws - 1, (int)floorf(fmaxf(u0, u1) - 0.5f + kEps) + 1);
          if (a <= b) { xlo = min(xlo, a); xhi = max(xhi, b); }
        }
      }
    }
    xlo = __reduce_min_sync(kFullMask, xlo);
    xhi = __reduce_max_sync(kFullMask, xhi);
    if (xlo <= xhi) {
      const int a8 = xlo & ~7, len8 = ((xhi + 8) & ~7) - a8;   // whole groups: [a8, a8 + len8) covers [xlo, xhi]
      const int r = y - wy0;
      if (lane == (r & 31)) {
        if (r < 32) { sxa[0] = a8; slen[0] = len8; } else { sxa[1] = a8; slen[1] = len8; }
      }
      total += len8;
      if (total > kRMax) return total;
    }
  }
  int i0 = slen[0], i1 = slen[1];                  // inclusive prefix sums over <= 64 rows
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int a = __shfl_up_sync(kFullMask, i0, o), b = __shfl_up_sync(kFullMask, i1, o);
    if (lane >= o) { i0 += a; i1 += b; }
  }
  const int tot0 = __shfl_sync(kFullMask, i0, 31), tot1 = __shfl_sync(kFullMask, i1, 31);
  if (lane < H) { RowSpan r; r.xa = sxa[0]; r.len = slen[0]; r.lenm1 = max(slen[0] - 1, 0); r.offp = i0 - slen[0] - sxa[0]; rows[lane] = r; }
  if (lane + 32 < H) { RowSpan r; r.xa = sxa[1]; r.len = slen[1]; r.lenm1 = max(slen[1] - 1, 0); r.offp = tot0 + i1 - slen[1] - sxa[1]; rows[lane + 32] = r; }
  ymin_out = wy0; H_out = H;
  __syncwarp();
  return tot0 + tot1;
}

// Inverse-depth interval of planes [s, s + n) (the schedule need not be monotone), by one converged warp.  False if not finite.
__device__ bool plane_interval(const float* __restrict__ invd, long long stride, int s, int n, float& dlo, float& dhi) {
  const int lane = threadIdx.x & 31;
  dlo = INFINITY; dhi = -INFINITY;
  float d[kMaxPlanes / 32];
#pragma unroll
  for (int k = 0; k < kMaxPlanes / 32; ++k) {
    const int i = lane + 32 * k;
    d[k] = (i < n) ? invd[(size_t)(s + i) * stride] : NAN;
  }
#pragma unroll
  for (int k = 0; k < kMaxPlanes / 32; ++k) {       // fminf / fmaxf drop nans: a nan plane has no valid sample
    dlo = fminf(dlo, d[k]); dhi = fmaxf(dhi, d[k]);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    dlo = fminf(dlo, __shfl_xor_sync(kFullMask, dlo, o));
    dhi = fmaxf(dhi, __shfl_xor_sync(kFullMask, dhi, o));
  }
  return (fabsf(dlo) < INFINITY) && (fabsf(dhi) < INFINITY);
}

// ---------------------------------------------------------------------------------------------------------
// warps 0 and 3: planning.  Warp 0 walks the work items and finds each chunk's band; warp 3 fills the per-plane table of the
// same descriptor at the same time (the two meet on a one-entry mailbox before the descriptor is published).
// ---------------------------------------------------------------------------------------------------------
struct PlanJob { int item, slot, s, len, stop; };   // plane-table job: planes [s, s + len) of `item` into descriptor `slot`

struct ItemGeo { int tile, n, v, hs, ws, by, bx; };
__device__ __forceinline__ ItemGeo item_geo(const PsweepArgs& A, int item, int tiles, int tiles_x) {
  ItemGeo g;
  const int q = item / tiles;
  g.tile = item - q * tiles;
  g.v = q / A.N;
  g.n = q - g.v * A.N;
  g.hs = A.hs[g.v]; g.ws = A.ws[g.v];
  g.by = g.tile / tiles_x; g.bx = g.tile - g.by * tiles_x;
  return g;
}

// inverse-depth schedule of sample n -> shared memory (planes [s_begin, s_end), at most kSchedMax); returns whether every entry
// is finite and the sequence monotone (the interval of a plane range is then its two end points)
__device__ bool load_schedule(const PsweepArgs& A, int n, float* sched) {
  const int lane = threadIdx.x & 31, S_item = A.s_end - A.s_begin;
  const float* invd = A.invd + (size_t)n * A.ist[0];
  __syncwarp();
  for (int i = lane; i < S_item; i += 32) sched[i] = invd[(size_t)(A.s_begin + i) * A.ist[1]];
  __syncwarp();
  bool asc = true, desc = true, fin = true;
  for (int i = lane; i < S_item; i += 32) {
    const float a = sched[i], b = sched[min(i + 1, S_item - 1)];
    asc = asc && (a <= b); desc = desc && (a >= b); fin = fin && (fabsf(a) < INFINITY);
  }
  return __all_sync(kFullMask, fin) && (__all_sync(kFullMask, asc) || __all_sync(kFullMask, desc));
}

__device__ void planner(const PsweepArgs& A, const __nv_bfloat16* key_split, ChunkDesc* descs, unsigned bars, int tw, int th, int tiles_x,
                        int tiles, int items, unsigned* counter, float* sched, PlanJob* job, int item_stride, int item_offset) {
  const int lane = threadIdx.x & 31;
  const int P = A.h * A.w, C = A.C;
  const unsigned bar_cready = bars + 8 * kBarCReady, bar_cfree = bars + 8 * kBarCFree, bar_jready = bars + 8 * kBarJobReady,
                 bar_jdone = bars + 8 * kBarJobDone;
  const int S_item = A.s_end - A.s_begin;
  const bool cached = S_item <= kSchedMax;
  unsigned seq = 0;
#ifdef RMVD_WS_STATS
  const long long pl_t0 = clock64();
#endif
  int prev_chunks = 1;
  Epi E{};
  int cur_n = -1, cur_v = -1;
  bool sched_mono = false;
  // Work items are handed out in index order through a global counter (the first gridDim.x statically): tiles differ in cost
  // by a factor of three, and neighbouring CTAs keep working on neighbouring tiles (shared source rows in L2).
  // `items` counts this launch's items; launch-local index j stands for item j * item_stride + item_offset (the ranks of a
  // tile-partitioned depth-slab group take every item_stride-th item each)
  int jitem = blockIdx.x;
  while (jitem < items) {
    const int item = jitem * item_stride + item_offset;
    unsigned nxt = 0;
    if (lane == 0) nxt = atomicAdd(counter, 1u);      // fetched now, needed after this item's chunks
    const ItemGeo g = item_geo(A, item, tiles, tiles_x);
    const int hs = g.hs, ws = g.ws;
    auto pixel_of_slot = [&](int slot) {
      const int ty = slot / tw, tx = slot - ty * tw;
      const int y = g.by * th + ty, x = g.bx * tw + tx;
      return (ty < th && y < A.h && x < A.w) ? y * A.w + x : -1;
    };
    const int p0 = pixel_of_slot(lane), p1 = pixel_of_slot(lane + 32);
    if (g.n != cur_n || g.v != cur_v) {                // calibration of this (sample, view): once per run of items
      E = epi_from_calib(A.Kk + g.n * 9, A.Ks[g.v] + g.n * 9, A.T[g.v] + g.n * 16, A.h, A.w, hs, ws);
      cur_v = g.v;
    }
    if (cached && g.n != cur_n && (cur_n < 0 || A.ist[0] != 0)) sched_mono = load_schedule(A, g.n, sched);   // warp 3 is idle here
    cur_n = g.n;
    const float* invd = A.invd + (size_t)g.n * A.ist[0];
    int s = A.s_begin, nchunks = 0;
    int len_hint = (S_item + max(prev_chunks - 1, 1) - 1) / max(prev_chunks - 1, 1);   // one chunk fewer than the last tile needed
    while (s < A.s_end) {
      const unsigned slot = seq % kDescs;
      if (seq >= kDescs) {
        WS_T0();
        mbar_wait(bar_cfree + 8 * slot, ((seq / kDescs) - 1) & 1);
        if (lane == 0) WS_ADD(0);
      }
      ChunkDesc* D = descs + slot;
      const int rem = A.s_end - s, L0 = min(rem, kMaxPlanes);
      int len = min(max(len_hint, 1), L0), R = -1, ymin = 0, H = 0;
      if (lane == 0) { job->item = item; job->slot = (int)slot; job->s = s; job->len = len; job->stop = 0; mbar_arrive(bar_jready); }
      D->pix[lane] = p0; D->pix[lane + 32] = p1;
      __syncwarp();
      bool direct = false;
      for (;;) {
        float dlo, dhi;
        WS_T0();
        bool finite;
        if (cached && sched_mono) {
          const float da = sched[s - A.s_begin], db = sched[s - A.s_begin + len - 1];
          dlo = fminf(da, db); dhi = fmaxf(da, db);
          finite = true;
        } else if (cached) {
          finite = plane_interval(sched, 1, s - A.s_begin, len, dlo, dhi);
        } else {
          finite = plane_interval(invd, A.ist[1], s, len, dlo, dhi);
        }
        R = finite ? band_warp(D->rows, D->pix, E, A.w, dlo, dhi, hs, ws, ymin, H) : -1;
        if (lane == 0) { WS_ADD(2); WS_CNT(1, 1); }
        if (R >= 0 && R <= kRMax) break;
        if (len <= 4) { direct = true; break; }
        int k = (rem + len - 1) / len;               // balanced split of the remaining planes into more chunks
        k = (R < 0) ? 2 * k : (k < 8 ? k + 1 : k + (k >> 2));
        len = max(4, min(len - 1, (rem + k - 1) / k));
      }
      if (lane == 0) {
        const int wg = (ws + 7) >> 3, G = hs * wg;
        D->E = E;
        D->key = key_split + ((size_t)g.n * tiles + g.tile) * (size_t)C * 128;
        D->src = reinterpret_cast<const __nv_bfloat16*>(A.src[g.v]) + (size_t)g.n * (C >> 5) * G * 512;
        D->out_base = (unsigned long long)(((size_t)(A.v0 + g.v) * A.N + g.n) * A.out_S) * (unsigned long long)P;
        D->hs = hs; D->ws = ws; D->wg = wg; D->G = G;
        D->n = g.n; D->v = g.v;
        D->kind = direct ? kDirect : kGemm;
        D->first = (s == A.s_begin) ? 1 : 0;
        D->last = (s + len >= A.s_end) ? 1 : 0;
        D->s0 = s; D->len = len; D->R = direct ? 0 : R; D->ymin = ymin; D->H = direct ? 0 : H;
        D->out_plane0 = A.out_s0 + (s - A.s_begin);
#ifdef RMVD_WS_STATS
        WS_CNT(5, 1); WS_CNT(6, R > 0 ? R : 0); WS_CNT(7, len);
#endif
      }
      {
        WS_T0();
        mbar_wait(bar_jdone, seq & 1);                 // the plane table of this descriptor is complete
        if (lane == 0) WS_ADD(3);
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_cready + 8 * slot);
      ++seq;
      s += len;
      ++nchunks;
      const int rem2 = A.s_end - s;
      if (rem2 > 0) { const int k2 = (rem2 + len - 1) / len; len_hint = (rem2 + k2 - 1) / k2; }
    }
    prev_chunks = nchunks;
    jitem = (int)__shfl_sync(kFullMask, nxt, 0) + (int)gridDim.x;
  }
#ifdef RMVD_WS_STATS
  if (lane == 0) atomicAdd(&g_ws_stats[4], (unsigned long long)(clock64() - pl_t0));
#endif
  const unsigned slot = seq % kDescs;
  if (seq >= kDescs) mbar_wait(bar_cfree + 8 * slot, ((seq / kDescs) - 1) & 1);
  if (lane == 0) { descs[slot].kind = kStop; job->stop = 1; mbar_arrive(bar_jready); }
  __syncwarp();
  if (lane == 0) mbar_arrive(bar_cready + 8 * slot);
}

// warp 3: per-plane terms of the coordinate chain, rounded like the reference, and the dead-plane bits: no pixel of the tile can
// have a valid sample on a dead plane (1/d <= 0, :499, or the plane's homography sends all four tile corners -- hence, with the
// denominator of one sign over the tile, the whole tile -- beyond the same edge of the source map by more than a pixel); the
// denominators must dominate the rounding errors of the per-pixel terms by a wide margin.  The job covers the longest chunk the
// planner may settle on; planes beyond the final length are simply not used.
__device__ void plane_tabler(const PsweepArgs& A, ChunkDesc* descs, unsigned bars, int tw, int th, int tiles_x, int tiles, const float* sched,
                             const PlanJob* job) {
  const int lane = threadIdx.x & 31;
  const unsigned bar_jready = bars + 8 * kBarJobReady, bar_jdone = bars + 8 * kBarJobDone;
  const bool cached = (A.s_end - A.s_begin) <= kSchedMax;
  Epi E{};
  float mag_u = 0.f, mag_v = 0.f, mag_k = 0.f;
  EpiPix cpix[4] = {};
  int cur_n = -1, cur_v = -1, cur_item = -1, hs = 1, ws = 1;
  const float* invd = A.invd;
  for (unsigned seq = 0;; ++seq) {
    mbar_wait(bar_jready, seq & 1);
    if (job->stop) break;
    const int item = job->item, s = job->s, len = job->len;
    ChunkDesc* D = descs + job->slot;
#ifdef RMVD_WS_STATS
    const long long pt_t0 = clock64();
#endif
    if (item != cur_item) {
      const ItemGeo g = item_geo(A, item, tiles, tiles_x);
      hs = g.hs; ws = g.ws;
      if (g.n != cur_n || g.v != cur_v) {
        E = epi_from_calib(A.Kk + g.n * 9, A.Ks[g.v] + g.n * 9, A.T[g.v] + g.n * 16, A.h, A.w, hs, ws);
        // magnitudes of the summands of u_h, v_h, k_h over the key map (rounding-error scale of the per-pixel terms)
        mag_u = fabsf(E.a) * (float)A.w + fabsf(E.b) * (float)A.h + fabsf(E.c);
        mag_v = fabsf(E.f) * (float)A.w + fabsf(E.g) * (float)A.h + fabsf(E.h);
        mag_k = fabsf(E.j) * (float)A.w + fabsf(E.k) * (float)A.h + fabsf(E.l);
        cur_n = g.n; cur_v = g.v;
        invd = A.invd + (size_t)g.n * A.ist[0];
      }
      const int xa = g.bx * tw, xb = min(xa + tw, A.w) - 1, ya = g.by * th, yb = min(ya + th, A.h) - 1;
      cpix[0] = epi_pixel(E, xa, ya); cpix[1] = epi_pixel(E, xb, ya);      // the four corner pixels of the tile
      cpix[2] = epi_pixel(E, xa, yb); cpix[3] = epi_pixel(E, xb, yb);
      cur_item = item;
    }
    const float edge_x = (float)ws + 1.0f, edge_y = (float)hs + 1.0f, span = (float)max(ws, hs) + 5.0f;
#pragma unroll 1
    for (int k8 = 0; 32 * k8 < len; ++k8) {
      const int i = lane + 32 * k8;
      const float d = (i < len) ? (cached ? sched[s - A.s_begin + i] : invd[(size_t)(s + i) * A.ist[1]]) : 1.0f;
      PlaneTerm pt;
      pt.ed = fmul(E.e, d); pt.id = fmul(E.i, d); pt.md = fmul(E.m, d); pt.zs = fdiv(1.0f, d);
      bool dead = !(pt.zs > 0.0f);
      if (dead) pt.zs = NAN;                           // never visible in the key view (:499): every compare with it is false
      if (i < len) D->planes[i] = pt;
      if (!dead) {
        const float den_min = 9.5367431640625e-7f *    // 16 * 2^-24
            (fmaxf(mag_u + fabsf(pt.ed), mag_v + fabsf(pt.id)) + span * (mag_k + fabsf(pt.md)));
        int npos = 0, nneg = 0, xl = 0, xr = 0, yt = 0, yb = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const float den = cpix[k].k_h + pt.md;
          const float aden = fabsf(den);
          const bool okd = (aden >= den_min) && (aden <= kDivHi);
          const float r = rcp_refined(den);
          const float ix = __fmaf_rn(cpix[k].u_h + pt.ed, r, -0.5f), iy = __fmaf_rn(cpix[k].v_h + pt.id, r, -0.5f);
          npos += (okd && den > 0.f); nneg += (okd && den < 0.f);
          xl += (ix < -3.0f); xr += (ix > edge_x); yt += (iy < -3.0f); yb += (iy > edge_y);
        }
        dead = (npos == 4 || nneg == 4) && (xl == 4 || xr == 4 || yt == 4 || yb == 4);
      }
      const unsigned bits = __ballot_sync(kFullMask, dead && i < len && A.dbg_ix0 == nullptr);
      if (lane == 0) D->dead[k8] = bits;
    }
#ifdef RMVD_WS_STATS
    if (lane == 0) atomicAdd(&g_ws_stats[15], (unsigned long long)(clock64() - pt_t0));
#endif
    __syncwarp();
    if (lane == 0) mbar_arrive(bar_jdone);
  }
}

// ---------------------------------------------------------------------------------------------------------
// warp 1: producer.  Lane l owns band rows l and l + 32 of the chunk; per stage it intersects them with the pass.
// ---------------------------------------------------------------------------------------------------------
__device__ void producer(const ChunkDesc* descs, unsigned bars, unsigned stage0, int C) {
  const int lane = threadIdx.x & 31;
  const int nkc = C / kKC;
  const unsigned bar_cready = bars + 8 * kBarCReady, bar_cfree = bars + 8 * kBarCFree, bar_full = bars + 8 * kBarFull,
                 bar_free = bars + 8 * kBarFree;
  unsigned g = 0;
  for (unsigned seq = 0;; ++seq) {
    const unsigned slot = seq % kDescs;
    {
      WS_T0();
      mbar_wait(bar_cready + 8 * slot, (seq / kDescs) & 1);
      if (lane == 0) WS_ADD(8);
    }
    const ChunkDesc* D = descs + slot;
    const int kind = D->kind;
    if (kind == kStop) break;
    const int R = D->R;
    if (D->first) {
      // the key tile of this item: C / 64 stages of [4 K steps][128 A' rows][32 bytes], copied to TMEM by the MMA thread
      const unsigned char* ksrc = reinterpret_cast<const unsigned char*>(D->key);
      for (int i = 0; i < (C + 63) / 64; ++i, ++g) {
        const unsigned sl = g % kStages;
        if (g >= kStages) mbar_wait(bar_free + 8 * sl, ((g / kStages) - 1) & 1);
        if (lane == 0) {
          const unsigned bytes = (unsigned)min(4, C / 16 - 4 * i) * 4096u, full = bar_full + 8 * sl;
          asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(full), "r"(bytes) : "memory");
          bulk_g2s(stage0 + sl * kStageBytes, ksrc + (size_t)i * kStageBytes, bytes, full);
        }
      }
    }
    if (kind == kGemm && R > 0) {
      const int H = D->H, ymin = D->ymin, wg = D->wg;
      int cs[2], n8[2], gofs[2];
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const int h = lane + 32 * j;
        cs[j] = 0; n8[j] = 0; gofs[j] = 0;
        if (h < H) {
          const RowSpan r = D->rows[h];
          cs[j] = r.offp + r.xa; n8[j] = r.len; gofs[j] = (ymin + h) * wg + (r.xa >> 3);
        }
      }
      const unsigned char* src = reinterpret_cast<const unsigned char*>(D->src);
      const size_t kcbytes = (size_t)D->G * 1024;
      const int npass = (R + kNT - 1) / kNT;
      for (int pass = 0; pass < npass; ++pass) {
        const int lo = pass * kNT, hi = min(R, lo + kNT);
        for (int kc = 0; kc < nkc; ++kc, ++g) {
          const unsigned sl = g % kStages;
          if (g >= kStages) {
            WS_T0();
            mbar_wait(bar_free + 8 * sl, ((g / kStages) - 1) & 1);   // the MMAs that read this slot are done
            if (lane == 0) WS_ADD(9);
          }
          const unsigned full = bar_full + 8 * sl;
          if (lane == 0) asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(full), "r"((unsigned)(hi - lo) * 128u) : "memory");
          const unsigned dst0 = stage0 + sl * kStageBytes;
          const unsigned char* srck = src + (size_t)kc * kcbytes;
#pragma unroll
          for (int j = 0; j < 2; ++j) {
            const int a = max(cs[j], lo), b = min(cs[j] + n8[j], hi);
            if (a < b) bulk_g2s(dst0 + (unsigned)(a - lo) * 128u, srck + ((size_t)(gofs[j] + ((a - cs[j]) >> 3)) << 10), (unsigned)(b - a) * 128u, full);
          }
        }
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(bar_cfree + 8 * slot);
  }
}

// ---------------------------------------------------------------------------------------------------------
// warp 2: MMA issuer (one elected lane).  Every chunk has at least one pass on the accumulator ring (an empty one for
// chunks without a band), so that this warp never falls behind the consumers.
// ---------------------------------------------------------------------------------------------------------
__device__ void mma_issuer(const ChunkDesc* descs, unsigned bars, unsigned stage0, unsigned tmem, int C) {
  const int nkc = C / kKC;
  const unsigned bar_cready = bars + 8 * kBarCReady, bar_full = bars + 8 * kBarFull, bar_free = bars + 8 * kBarFree,
                 bar_acc = bars + 8 * kBarAcc, bar_drained = bars + 8 * kBarDrained, bar_afree = bars + 8 * kBarAFree;
  const unsigned long long desc0 = umma_desc(stage0, 128, 1024);
  const unsigned idesc0 = umma_idesc_bf16(2 * kPx, 0);
  unsigned g = 0, ap = 0, af_seq = 0;
  bool mma_pending = false;                          // MMAs issued since the last key copy
  for (unsigned seq = 0;; ++seq) {
    const unsigned slot = seq % kDescs;
    {
      WS_T0();
      mbar_wait(bar_cready + 8 * slot, (seq / kDescs) & 1);
      WS_ADD(10);
    }
    const ChunkDesc* D = descs + slot;
    const int kind = D->kind;
    if (kind == kStop) break;
    if (D->first) {
      // Key tile -> TMEM (A' columns) by tcgen05.cp, one 128 x 256-bit copy per K = 16 step.  The copies execute in issue order with
      // the MMAs that follow; the MMAs of the previous item that still read A' are waited for explicitly.
      if (mma_pending) {
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n" ::"r"(bar_afree) : "memory");
        WS_T0();
        mbar_wait(bar_afree, af_seq & 1);
        WS_ADD(11);
        ++af_seq;
        mma_pending = false;
      }
      for (int i = 0; i < (C + 63) / 64; ++i, ++g) {
        const unsigned sl = g % kStages;
        mbar_wait(bar_full + 8 * sl, (g / kStages) & 1);
        asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory");
        const int nks = min(4, C / 16 - 4 * i);
        for (int k = 0; k < nks; ++k) {
          const unsigned long long kd = umma_desc(stage0 + sl * kStageBytes + k * 4096, 128, 256);
          asm volatile("tcgen05.cp.cta_group::1.128x256b [%0], %1;\n" ::"r"(tmem + kTmemA + (4 * i + k) * 8), "l"(kd) : "memory");
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n" ::"r"(bar_free + 8 * sl) : "memory");
      }
    }
    const int R = (kind == kGemm) ? D->R : 0;
    const int npass = (R + kNT - 1) / kNT;
    if (npass == 0) {
      const unsigned buf = ap % kNBuf, use = ap / kNBuf;
      if (use >= 1) mbar_wait(bar_drained + 8 * buf, (use - 1) & 1);
      mbar_arrive(bar_acc + 8 * buf);
      ++ap;
      continue;
    }
    for (int pass = 0; pass < npass; ++pass, ++ap) {
      const int ncol = min(R - pass * kNT, kNT);
      const unsigned idesc = idesc0 | ((unsigned)(((ncol + 15) & ~15) >> 3) << 17);
      const unsigned buf = ap % kNBuf, use = ap / kNBuf;
      if (use >= 1) {
        WS_T0();
        mbar_wait(bar_drained + 8 * buf, (use - 1) & 1);   // this accumulator's previous contents have been read out
        WS_ADD(12);
      }
      asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory");
      const unsigned dcol = tmem + buf * kNT;
      for (int kc = 0; kc < nkc; ++kc, ++g) {
        const unsigned sl = g % kStages;
        {
          WS_T0();
          mbar_wait(bar_full + 8 * sl, (g / kStages) & 1);
          WS_ADD(13);
          WS_CNT(14, 1);
        }
        asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory");
        // descriptors differ only in their start-address field (units of 16 bytes): stage base, then 256 bytes per K = 16 step
        // (two 16-byte chunks); the hi steps are chunks 0..3, the lo steps chunks 4..7 and re-use the A' slices of the hi steps
        const unsigned long long d0 = desc0 + (unsigned long long)(sl * (kStageBytes >> 4));
        const unsigned a_addr = tmem + kTmemA + kc * (kKC / 2);
        tc_mma(dcol, a_addr, d0, idesc, kc > 0 ? 1u : 0u);
        tc_mma(dcol, a_addr + 8, d0 + 16, idesc, 1u);
        tc_mma(dcol, a_addr, d0 + 32, idesc, 1u);
        tc_mma(dcol, a_addr + 8, d0 + 48, idesc, 1u);
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n" ::"r"(bar_free + 8 * sl) : "memory");
      }
      asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n" ::"r"(bar_acc + 8 * buf) : "memory");
      mma_pending = true;
    }
  }
}

// ---------------------------------------------------------------------------------------------------------
// consumers
// ---------------------------------------------------------------------------------------------------------
// dot product of one key pixel with one source texel from the split copies (slow paths only)
__device__ float dot_thread(const ChunkDesc* D, int C, int slot, int y, int x) {
  float acc = 0.f;
  for (int c = 0; c < C; ++c) acc = fmaf(key_at(D->key, slot, c), src_at(D, y, x, c), acc);
  return acc;
}

// One sample through the exact chain, evaluated and stored (about one sample in 10^3 of the filtered locate below).
template <int kMode>
__device__ __noinline__ void sample_slow(const ChunkDesc* D, int p, int pl, int slot) {
  const WsConst& K = s_k;
  const int hs = D->hs, ws = D->ws, R = D->R, C = K.C;
  const int ymin = D->ymin, H = D->H, Hm1 = max(H - 1, 0), Rm1 = max(R - 1, 0);
  const float* g = K.G + slot * kGPitch;
  const Pixel px = pixel_consts(D->E, p, K.w);
  Sample L;
  locate(px, D->planes[pl], hs, ws, L);
  float val = 0.f;
  if (L.valid) {
    // Corners outside the map carry weight exactly 0, so they are read from a clamped (finite) column instead of being
    // branched around; a corner WITH weight outside the predicted spans is a miss and is evaluated directly.
    const int h0 = L.y0 - ymin;
    const bool rowA = (unsigned)h0 < (unsigned)H, rowB = (unsigned)(h0 + 1) < (unsigned)H;
    const RowSpan rA = D->rows[min(max(h0, 0), Hm1)], rB = D->rows[min(max(h0 + 1, 0), Hm1)];
    const int xl = L.x0, xr = L.x0 + 1;
    auto in_span = [](const RowSpan& r, int x) { return (unsigned)(x - r.xa) < (unsigned)r.len; };
    const bool inA_l = rowA && in_span(rA, xl), inA_r = rowA && in_span(rA, xr);
    const bool inB_l = rowB && in_span(rB, xl), inB_r = rowB && in_span(rB, xr);
    const bool miss = (R == 0) || (L.nw != 0.f && !inA_l) || (L.ne != 0.f && !inA_r) || (L.sw != 0.f && !inB_l) || (L.se != 0.f && !inB_r);
    float g_nw = 0.f, g_ne = 0.f, g_sw = 0.f, g_se = 0.f;
    if (!miss) {
      const int cA = rA.offp + xl, cB = rB.offp + xl;
      g_nw = g[min(max(cA, 0), Rm1)]; g_ne = g[min(max(cA + 1, 0), Rm1)];
      g_sw = g[min(max(cB, 0), Rm1)]; g_se = g[min(max(cB + 1, 0), Rm1)];
    } else {
      const int x0c = clampi(L.x0, 0, ws - 1), x1c = clampi(L.x0 + 1, 0, ws - 1);
      const int y0c = clampi(L.y0, 0, hs - 1), y1c = clampi(L.y0 + 1, 0, hs - 1);
      g_nw = (L.nw != 0.f) ? dot_thread(D, C, slot, y0c, x0c) : 0.f;
      g_ne = (L.ne != 0.f) ? dot_thread(D, C, slot, y0c, x1c) : 0.f;
      g_sw = (L.sw != 0.f) ? dot_thread(D, C, slot, y1c, x0c) : 0.f;
      g_se = (L.se != 0.f) ? dot_thread(D, C, slot, y1c, x1c) : 0.f;
    }
    val = (g_nw * L.nw + g_ne * L.ne + g_sw * L.sw + g_se * L.se) * K.inv_scale;
  }
  const size_t o = (size_t)D->out_base + (size_t)(D->out_plane0 + pl) * K.P + p;
  store_out<kMode>(K, K.corr, K.mask, (size_t)0, o, val, L.valid ? 1.f : 0.f);
  if (kMode == kModeDebug) {
    K.dbg_ix0[o] = L.x0;
    K.dbg_iy0[o] = L.y0;
    if (K.dbg_geom) K.dbg_geom[o] = L.geom ? 1 : 0;
  }
}

// Filtered locate.  Only the INTEGER corner (floor of the un-normalised coordinate) and the 0/1 mask have to be the
// reference's bit for bit; the bilinear weights only to 1e-4.  With U = nu / den the reference's chain
// (us = nu/den; xg = 2 us/ws - 1; ix = fma(xg + 1, ws/2, -0.5), five roundings) stays within u (5.5 ws + 10.5) of
// U - 0.5 (u = 2^-24, |U| <= ws + 2), and ix' = fma(nu, rcp(den), -0.5) with an exactly rounded reciprocal within
// u (3 ws + 6.5): whenever ix' is further than eps = u (16 ws + 64) from an integer, floor(ix') IS the reference's floor.
// The reciprocal here is MUFU's (relative error <= 2^-23 = 2u, i.e. <= 2u (ws + 2) more on ix'): the margin
// u (10.5 ws + 53.5) left by the two bounds covers it several times (tools/filter_check.py perturbs the reciprocal by
// +-2 ulp).  A sample is settled by the cheap chain when both floors are trusted (or the coordinate is beyond
// [-3, size + 1], where the saturated index is decided) and it is either interior (all four corners inside the map: the
// in-range weight sum is 1 +- 6u >= 0.9999, mask = geometric mask) or has a corner column / row outside the map with the
// in-range weight below 0.9999 by more than eps (mask = 0).  Everything else -- about one sample in 10^3 -- takes
// sample_slow(), the exact chain.  tests/test_gpu_planesweep.py compares indices, geometric masks and masks bit for bit.
struct FastConst {
  float u_h, v_h, k_h;
  float sg, zp;           // visibility in the source as one compare: zs * sg > zp (planesweep_corr.py:500-504)
  float tin_x, tin_y;     // 0.5 - trust margin of an interior floor: trusted iff |frac - 0.5| < this
  float tout_x, tout_y;   // the same where a corner column / row is outside the map (margin covers the 0.9999 threshold)
  int ymin, Hm1;
  unsigned wsm1, hsm1;
};
struct FastOut { float val; int xi, yi; bool valid, geom, settled; };

__device__ __forceinline__ FastOut sample_fast(const RowSpan* __restrict__ rows, const FastConst& K, const float* __restrict__ g,
                                               const PlaneTerm pt, float inv_scale) {
  FastOut o;
  const float den = fadd(K.k_h, pt.md), nu = fadd(K.u_h, pt.ed), nv = fadd(K.v_h, pt.id);      // :334, :343
  o.geom = pt.zs * K.sg > K.zp;                                  // :499-506 (zs is NaN where 1/d <= 0)
  const bool den_ok = fabsf(den) <= kDivHi;
  float r;                                                        // 1 ulp reciprocal: inside the trust margins (see above)
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(den));
  const float ix = __fmaf_rn(nu, r, -0.5f), iy = __fmaf_rn(nv, r, -0.5f);
  const float fx = ix - floorf(ix), fy = iy - floorf(iy);        // exact
  o.xi = __float2int_rd(ix); o.yi = __float2int_rd(iy);          // saturating; nan -> 0
  const bool int_x = (unsigned)o.xi < K.wsm1, int_y = (unsigned)o.yi < K.hsm1;
  const bool trust_x = fabsf(fx - 0.5f) < (int_x ? K.tin_x : K.tout_x);      // frac further than the margin from 0 and 1; false for nan
  const bool trust_y = fabsf(fy - 0.5f) < (int_y ? K.tin_y : K.tout_y);
  o.valid = o.geom && int_x && int_y;
  const int h0 = o.yi - K.ymin;
  const bool rows_ok = (unsigned)h0 < (unsigned)K.Hm1;
  const int hc = rows_ok ? h0 : 0;
  const RowSpan rA = rows[hc], rB = rows[hc + 1];
  const bool in_band = rows_ok & ((unsigned)(o.xi - rA.xa) < (unsigned)rA.lenm1) & ((unsigned)(o.xi - rB.xa) < (unsigned)rB.lenm1);
  const bool use = o.valid && in_band;
  const float* ga = g + (use ? rA.offp + o.xi : 0);              // unused samples read columns 0, 1 (discarded)
  const float* gb = g + (use ? rB.offp + o.xi : 0);
  const float a0 = ga[0], a1 = ga[1], b0 = gb[0], b1 = gb[1];
  const float top = fmaf(fx, a1 - a0, a0), bot = fmaf(fx, b1 - b0, b0);
  o.val = use ? fmaf(fy, bot - top, top) * inv_scale : 0.f;
  o.settled = den_ok && trust_x && trust_y && (in_band || !o.valid);
  return o;
}

// warp = 32 adjacent tile slots x one plane per step (one 128-byte row of corr and of mask), two planes in flight per thread
template <int kMode>
__device__ __forceinline__ void sample_chunk(const ChunkDesc* D, const float* G, const Pixel& px, int p, int cw, int lane) {
  if (p < 0) return;
  const WsConst& W = s_k;
  const int hs = D->hs, ws = D->ws, len = D->len;
  const float inv_scale = W.inv_scale;
  constexpr float kU = 5.9604645e-8f;   // 2^-24
  FastConst K;
  K.u_h = px.u_h; K.v_h = px.v_h; K.k_h = px.k_h;
  K.sg = (px.mode == 2) ? -1.0f : 1.0f;
  K.zp = (px.mode == 1) ? px.z_pole : (px.mode == 2) ? -px.z_pole : (px.mode == 3) ? -INFINITY : INFINITY;
  {
    const float eps_x = kU * (16.0f * (float)ws + 64.0f), eps_y = kU * (16.0f * (float)hs + 64.0f);
    K.tin_x = 0.5f - eps_x; K.tin_y = 0.5f - eps_y;
    K.tout_x = K.tin_x - 1.2e-4f; K.tout_y = K.tin_y - 1.2e-4f;
  }
  K.ymin = D->ymin; K.Hm1 = max(D->H - 1, 0);
  K.wsm1 = (unsigned)(ws - 1); K.hsm1 = (unsigned)(hs - 1);
  const int slot = (cw & 1) * 32 + lane;
  const float* g = G + slot * kGPitch;
  const RowSpan* rows = D->rows;
  const PlaneTerm* planes = D->planes;
  // outputs of this chunk: 64-bit bases (plane 0 of the chunk, this thread's pixel) + 32-bit element offsets
  const size_t base = (size_t)D->out_base + (size_t)D->out_plane0 * W.P + p;
  float* const corr = W.corr + base; float* const mask = W.mask + base;
  const unsigned P = (unsigned)W.P, step = (unsigned)kStep * P;
  auto emit = [&](unsigned oo, int pl, const FastOut& r) {
    if (r.settled) {
      store_out<kMode>(W, corr, mask, base, oo, r.val, r.valid ? 1.f : 0.f);
      if (kMode == kModeDebug) {
        W.dbg_ix0[base + oo] = min(max(r.xi, -2), ws);   // sat_index()
        W.dbg_iy0[base + oo] = min(max(r.yi, -2), hs);
        if (W.dbg_geom) W.dbg_geom[base + oo] = r.geom ? 1 : 0;
      }
    } else {
      sample_slow<kMode>(D, p, pl, slot);
    }
  };
  const unsigned* dead = D->dead;
  auto is_dead = [&](int q) { return (dead[q >> 5] >> (q & 31)) & 1u; };
  int pl = cw >> 1;
  unsigned o = (unsigned)pl * P;
#ifndef WS_SU
#define WS_SU 2
#endif
  constexpr int kSU = WS_SU;                             // samples in flight per thread
  for (; pl + (kSU - 1) * kStep < len; pl += kSU * kStep, o += kSU * step) {
    bool all_dead = true;
#pragma unroll
    for (int j = 0; j < kSU; ++j) all_dead = all_dead && is_dead(pl + j * kStep);
    if (all_dead) {                                      // warp-uniform
#pragma unroll
      for (int j = 0; j < kSU; ++j) store_out<kMode>(W, corr, mask, base, o + j * step, 0.f, 0.f);
      continue;
    }
    FastOut r[kSU];
#pragma unroll
    for (int j = 0; j < kSU; ++j) r[j] = sample_fast(rows, K, g, planes[pl + j * kStep], inv_scale);
#pragma unroll
    for (int j = 0; j < kSU; ++j) emit(o + j * step, pl + j * kStep, r[j]);
  }
  for (; pl < len; pl += kStep, o += step) emit(o, pl, sample_fast(rows, K, g, planes[pl], inv_scale));
}

// Irregular chunk: every sample located with the reference chain of common.cuh and evaluated directly, the warp sharing each
// valid sample's channel reduction.
template <int kMode>
__device__ void direct_chunk(const PsweepArgs& A, const ChunkDesc* D, int cw, int lane) {
  const WsConst& W = s_k;
  const int C = W.C, P = W.P, hs = D->hs, ws = D->ws, len = D->len, n = D->n;
  const int slot0 = (cw & 1) * 32;
  const int p = D->pix[slot0 + lane];
  for (int pl = cw >> 1; pl < len; pl += kStep) {
    const int s = D->s0 + pl;
    bool valid = false, geom = false;
    float wnw = 0.f, wne = 0.f, wsw = 0.f, wse = 0.f;
    int x0 = 0, y0 = 0, x0c = 0, x1c = 0, y0c = 0, y1c = 0;
    if (p >= 0) {
      const int yi = p / A.w, xi = p - yi * A.w;
      const EpiPix ep = epi_pixel(D->E, xi, yi);
      const float d = A.invd[n * A.ist[0] + s * A.ist[1] + yi * A.ist[2] + xi * A.ist[3]];
      float us, vs;
      epi_sample(D->E, ep, d, us, vs, geom);
      const Bilinear b = bilinear_setup(pix_to_unnorm(us, ws), pix_to_unnorm(vs, hs), hs, ws);
      valid = sampling_mask(b.msum) && geom;
      wnw = b.nw; wne = b.ne; wsw = b.sw; wse = b.se;
      x0 = sat_index(b.x0f, ws); y0 = sat_index(b.y0f, hs);
      x0c = clampi(x0, 0, ws - 1); x1c = clampi(x0 + 1, 0, ws - 1);
      y0c = clampi(y0, 0, hs - 1); y1c = clampi(y0 + 1, 0, hs - 1);
    }
    unsigned vm = __ballot_sync(kFullMask, valid);
    float res = 0.f;
    while (vm) {
      const int j = __ffs(vm) - 1;
      vm &= vm - 1;
      const int qx0 = __shfl_sync(kFullMask, x0c, j), qx1 = __shfl_sync(kFullMask, x1c, j);
      const int qy0 = __shfl_sync(kFullMask, y0c, j), qy1 = __shfl_sync(kFullMask, y1c, j);
      const float nw = __shfl_sync(kFullMask, wnw, j), ne = __shfl_sync(kFullMask, wne, j);
      const float sw = __shfl_sync(kFullMask, wsw, j), se = __shfl_sync(kFullMask, wse, j);
      float a = 0.f;
      for (int c = lane; c < C; c += 32) {
        const float k = key_at(D->key, slot0 + j, c);
        a = fmaf(k, src_at(D, qy0, qx0, c) * nw + src_at(D, qy0, qx1, c) * ne + src_at(D, qy1, qx0, c) * sw + src_at(D, qy1, qx1, c) * se, a);
      }
      a = warp_sum(a);
      if (lane == j) res = a;
    }
    if (p >= 0) {
      const size_t o = (size_t)D->out_base + (size_t)(D->out_plane0 + pl) * P + p;
      store_out<kMode>(W, W.corr, W.mask, (size_t)0, o, valid ? res * W.inv_scale : 0.f, valid ? 1.f : 0.f);
      if (kMode == kModeDebug) {
        W.dbg_ix0[o] = x0;
        W.dbg_iy0[o] = y0;
        if (W.dbg_geom) W.dbg_geom[o] = geom ? 1 : 0;
      }
    }
  }
}

// push mode: one record per finished chunk {view-sample index, first output plane, planes, first pixel of the tile}; the slot's flag
// carries the call's epoch once the record AND the chunk's output rows are visible device-wide
__device__ __forceinline__ void publish_chunk(unsigned* q_head, unsigned q_cap, unsigned* tail, unsigned epoch, const int4& rec) {
  const unsigned slot = atomicAdd(tail, 1u);
  if (slot >= q_cap) return;                          // the host sized the queue for the worst case; never expected
  int4* recs = reinterpret_cast<int4*>(q_head + 4 + q_cap);
  recs[slot] = rec;
  __threadfence();                                    // cumulative: the consumer barrier before this call ordered every warp's stores before it
  asm volatile("st.release.gpu.global.u32 [%0], %1;\n" ::"l"(q_head + 4 + slot), "r"(epoch) : "memory");
}

template <int kMode>
__device__ void consumer(const PsweepArgs& A, ChunkDesc* descs, float* G, unsigned bars, unsigned tmem, unsigned* q_head, unsigned q_cap,
                         unsigned* q_tail, unsigned epoch) {
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int cw = warp - 4, q = warp & 3, part = (warp >> 2) - 1;   // TMEM lane quadrant, which of its four warps
  const unsigned bar_cready = bars + 8 * kBarCReady, bar_cfree = bars + 8 * kBarCFree, bar_acc = bars + 8 * kBarAcc,
                 bar_drained = bars + 8 * kBarDrained;
  Pixel px{};
  int p = -1;
  unsigned ap = 0;
  const int slot_px = 16 * q + (lane & 15);           // G row this lane drains
  const bool lower = lane < 16;
#ifdef RMVD_WS_STATS
  const bool rec = (t == 128);
  const long long cs_t0 = clock64();
#endif
  int4 prev_rec = make_int4(0, 0, 0, 0);
  bool have_prev = false;
  for (unsigned seq = 0;; ++seq) {
    const unsigned slot = seq % kDescs;
    {
      WS_T0();
      mbar_wait(bar_cready + 8 * slot, (seq / kDescs) & 1);
#ifdef RMVD_WS_STATS
      if (rec) WS_ADD(16);
#endif
    }
    const ChunkDesc* D = descs + slot;
    const int kind = D->kind;
    if (kind == kStop) break;
    if (D->first) {
      p = D->pix[(cw & 1) * 32 + lane];
      px = pixel_consts(D->E, max(p, 0), A.w);
#ifdef RMVD_WS_STATS
      if (rec) WS_CNT(24, 1);
#endif
    }
    const int R = (kind == kGemm) ? D->R : 0;
    const int npass = (R + kNT - 1) / kNT;
    {
      WS_T0();
      cons_bar();                                      // every warp has finished sampling the previous chunk: G may be overwritten
#ifdef RMVD_WS_STATS
      if (rec) WS_ADD(17);
#endif
    }
    if (kMode == kModePush && t == 128) {              // ... and its output rows are complete: hand the previous chunk to the pusher
      if (have_prev) publish_chunk(q_head, q_cap, q_tail, epoch, prev_rec);
      prev_rec = make_int4((int)(D->out_base / ((unsigned long long)A.out_S * (unsigned long long)(A.h * A.w))), D->out_plane0, D->len, D->pix[0]);
      have_prev = true;
    }
#ifdef RMVD_WS_STATS
    const long long dr_t0 = clock64();
    long long acc_wait = 0;
#endif
    for (int pass = 0; pass < max(npass, 1); ++pass, ++ap) {
      const unsigned buf = ap % kNBuf;
#ifdef RMVD_WS_STATS
      const long long aw0 = clock64();
#endif
      mbar_wait(bar_acc + 8 * buf, (ap / kNBuf) & 1);
#ifdef RMVD_WS_STATS
      acc_wait += clock64() - aw0;
#endif
      asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory");
      if (pass < npass) {
#pragma unroll 1
        for (int hb = 0; hb < kNT / 32; ++hb) {
          const int b = pass * (kNT / 32) + hb, col0 = b * 32;
          if (col0 < R && hb == part) {
            // Lane l < 16 holds the key_hi row of its pixel, lane l + 16 the key_lo row of the same pixel.  One exchange serves
            // two columns: the lower lane finishes column 2i, the upper lane column 2i + 1; both halves of the warp store.
            float a[32];
            tmem_ld32(tmem + buf * kNT + hb * 32 + ((unsigned)(q * 32) << 16), a);
            float* gcol = G + slot_px * kGPitch + col0 + (lower ? 0 : 1);
            const int left = R - col0 - (lower ? 0 : 1);
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              const float own = lower ? a[2 * i] : a[2 * i + 1];
              const float v = own + __shfl_xor_sync(kFullMask, lower ? a[2 * i + 1] : a[2 * i], 16);
              if (2 * i < left) gcol[2 * i] = v;
            }
          }
        }
      }
      asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory");
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_drained + 8 * buf);
    }
#ifdef RMVD_WS_STATS
    if (rec) { atomicAdd(&g_ws_stats[18], (unsigned long long)acc_wait); atomicAdd(&g_ws_stats[19], (unsigned long long)(clock64() - dr_t0 - acc_wait)); }
    const long long kl_t0 = clock64();
#endif
#ifdef RMVD_WS_STATS
    if (rec) atomicAdd(&g_ws_stats[20], (unsigned long long)(clock64() - kl_t0));
    const long long b2_t0 = clock64();
#endif
    cons_bar();                                        // G is complete
#ifdef RMVD_WS_STATS
    if (rec) atomicAdd(&g_ws_stats[21], (unsigned long long)(clock64() - b2_t0));
    const long long sm_t0 = clock64();
#endif
    if (kind == kGemm) sample_chunk<kMode>(D, G, px, p, cw, lane);
    else direct_chunk<kMode>(A, D, cw, lane);
    __syncwarp();
#ifdef RMVD_WS_STATS
    if (rec) atomicAdd(&g_ws_stats[22], (unsigned long long)(clock64() - sm_t0));
#endif
    if (lane == 0) mbar_arrive(bar_cfree + 8 * slot);
  }
  if (kMode == kModePush) {                            // the last chunk of this CTA
    cons_bar();
    if (t == 128 && have_prev) publish_chunk(q_head, q_cap, q_tail, epoch, prev_rec);
  }
#ifdef RMVD_WS_STATS
  if (rec) atomicAdd(&g_ws_stats[23], (unsigned long long)(clock64() - cs_t0));
#endif
}

template <int kMode>
__global__ void __launch_bounds__(kThreads, 1) psweep_ws_kernel(const PsweepArgs A, const __nv_bfloat16* __restrict__ key_split, int tw,
                                                                int th, int tiles_x, int tiles, int items, unsigned* counter,
                                                                const __grid_constant__ PeerArgs PA) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ unsigned s_tmem;
  unsigned char* stage = smem_raw + ((1024 - (smem_u32(smem_raw) & 1023)) & 1023);
  float* G = reinterpret_cast<float*>(stage + kStages * kStageBytes);
  ChunkDesc* descs = reinterpret_cast<ChunkDesc*>(G + kPx * kGPitch);
  unsigned long long* bar_mem = reinterpret_cast<unsigned long long*>(descs + kDescs);
  float* sched = reinterpret_cast<float*>(bar_mem + kNumBars);
  PlanJob* job = reinterpret_cast<PlanJob*>(sched + kSchedMax);
  const unsigned bars = smem_u32(bar_mem);
  const int t = threadIdx.x, warp = t >> 5;

  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(smem_u32(&s_tmem)), "r"(kTmemCols));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n");
  }
  if (t == 32) {
    for (int b = 0; b < kNumBars; ++b) {
      const int count = (b >= kBarDrained && b < kBarCReady) ? kConsWarps            // drained: every consumer warp
                        : (b >= kBarCFree && b < kBarAFree) ? kConsWarps + 1          // descriptor free: consumers + producer
                        : 1;
      asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(bars + 8 * b), "r"(count));
    }
    asm volatile("fence.mbarrier_init.release.cluster;\n");
    s_k.corr = A.corr; s_k.mask = A.mask; s_k.dbg_ix0 = A.dbg_ix0; s_k.dbg_iy0 = A.dbg_iy0; s_k.dbg_geom = A.dbg_geom;
    s_k.G = G; s_k.P = A.h * A.w; s_k.C = A.C; s_k.w = A.w;
    s_k.inv_scale = (A.norm == RMVD_NORM_DIM) ? 1.0f / sqrtf((float)A.C) : 1.0f;    // corr / sqrt(C), :178-179
    s_k.npeers = PA.n;
    for (int g = 0; g < kMaxPeers; ++g) { s_k.pcorr[g] = PA.corr[g]; s_k.pmask[g] = PA.mask[g]; }
    s_k.mc_corr = PA.n > 0 ? PA.corr_mc : nullptr; s_k.mc_mask = PA.n > 0 ? PA.mask_mc : nullptr;
  }
  asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory");
  const unsigned tmem = s_tmem;

  if (warp < 4) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;\n" ::"n"(kRegsControl));
    if (warp == 0) {
      planner(A, key_split, descs, bars, tw, th, tiles_x, tiles, items, counter, sched, job, PA.n > 0 ? PA.item_stride : 1, PA.n > 0 ? PA.item_offset : 0);
    } else if (warp == 1) {
      producer(descs, bars, smem_u32(stage), A.C);
    } else if (warp == 2) {
      if (elect_one()) mma_issuer(descs, bars, smem_u32(stage), tmem, A.C);
    } else {
      plane_tabler(A, descs, bars, tw, th, tiles_x, tiles, sched, job);
    }
  } else {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;\n" ::"n"(kRegsConsumer));
    consumer<kMode>(A, descs, G, bars, tmem, PA.q_head, PA.q_cap, counter + 1, PA.epoch);
  }

  if (kMode == kModePeers) __threadfence_system();                // this thread's peer stores are ordered before the handshake below
  asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" ::"r"(tmem), "r"(kTmemCols));
  if (kMode == kModePush && t == 0) {
    // the last CTA of the launch closes the queue: total number of records, then the call's epoch in the final flag
    __threadfence();
    const unsigned prev = atomicAdd(PA.done_ctas, 1u);
    if (prev == gridDim.x - 1) {
      *PA.done_ctas = 0u;
      const unsigned total = atomicAdd(counter + 1, 0u);
      PA.q_head[0] = total < PA.q_cap ? total : PA.q_cap;
      __threadfence();
      asm volatile("st.release.gpu.global.u32 [%0], %1;\n" ::"l"(PA.q_head + 1), "r"(PA.epoch) : "memory");
    }
  }
  if (kMode == kModePeers && t == 0) {
    // End-of-kernel handshake of the depth-slab mode: the last CTA of this rank publishes the call's epoch in every rank's flag
    // array and waits until all ranks have published theirs, so that when this kernel completes, every rank's rows have landed
    // in this rank's tensors (no separate barrier or collective launch).
    const unsigned prev = atomicAdd(PA.done_ctas, 1u);
    if (prev == gridDim.x - 1) {
      *PA.done_ctas = 0u;                              // next launch (stream order)
      __threadfence_system();
      for (int g = 0; g < PA.n; ++g) asm volatile("st.release.sys.global.u32 [%0], %1;\n" ::"l"(PA.flags[g] + PA.rank), "r"(PA.epoch) : "memory");
      for (int g = 0; g < PA.n; ++g) {
        unsigned have = 0, polls = 0;
        for (;;) {
          asm volatile("ld.acquire.sys.global.u32 %0, [%1];\n" : "=r"(have) : "l"(PA.flags[PA.rank] + g) : "memory");
          if ((int)(have - PA.epoch) >= 0) break;
          __nanosleep(200);
          if (++polls > (1u << 24)) __trap();         // a peer never arrived (several seconds): fail loudly rather than hang
        }
      }
    }
  }
}

static size_t smem_bytes() {
  return 1024 + (size_t)kStages * kStageBytes + sizeof(float) * kPx * kGPitch + sizeof(ChunkDesc) * kDescs + 8 * kNumBars +
         sizeof(float) * kSchedMax + sizeof(PlanJob);
}

}  // namespace ws

size_t psweep_ws_src_bytes(int N, int C, int hs, int ws) {
  return (size_t)N * C * hs * ((ws + 7) & ~7) * sizeof(float);
}

// Tile rectangle (tw x th <= 64 key pixels): fewest tiles wins, taller tiles weighted by the growth of their bands (one more
// source row per tile row: 32 x 2 costs ~5 % more per tile than 64 x 1, 16 x 4 ~20 %, 8 x 8 ~90 %).  Widths are rounded up to
// multiples of 8 pixels where that fits, so output rows of a warp start on 32-byte boundaries.
void psweep_ws_tile_shape(int h, int w, int* tw_out, int* th_out) {
  int tw = w, th = 1;
  double best = 1e30;
  for (int hh = 1; hh <= ws::kPx && hh <= h; ++hh) {
    const int wmax = ws::kPx / hh;
    const int nseg = (w + wmax - 1) / wmax;
    int wt = (w + nseg - 1) / nseg;
    if (((wt + 7) & ~7) <= wmax) wt = (wt + 7) & ~7;
    wt = wt < w ? wt : w;
    const double f = hh <= 3 ? 1.0 + 0.05 * (hh - 1) : (hh <= 8 ? 1.2 + 0.175 * (hh - 4) : 1.9 + 0.25 * (hh - 8));
    const double cost = f * ((w + wt - 1) / wt) * ((h + hh - 1) / hh);
    if (cost < best - 1e-9) { best = cost; tw = wt; th = hh; }
  }
  static const char* const ov = getenv("RMVD_WS_TILE");        // experiments: "tw,th"
  if (ov) {
    int a = 0, b = 0;
    if (sscanf(ov, "%d,%d", &a, &b) == 2 && a > 0 && b > 0 && a * b <= ws::kPx) { tw = a < w ? a : w; th = b < h ? b : h; }
  }
  *tw_out = tw; *th_out = th;
}

size_t psweep_ws_key_bytes(int N, int C, int h, int w) {
  int tw, th;
  psweep_ws_tile_shape(h, w, &tw, &th);
  const size_t tiles = (size_t)((w + tw - 1) / tw) * ((h + th - 1) / th);
  return (size_t)N * tiles * C * 256;
}

int repack_key_tiles(const float* key, void* dst, int N, int C, int h, int w, int l2norm, cudaStream_t st, bool act, float slope,
                     float* act_out) {
  int tw, th;
  psweep_ws_tile_shape(h, w, &tw, &th);
  const int tiles_x = (w + tw - 1) / tw, tiles = tiles_x * ((h + th - 1) / th);
  if (N > 65535) return (int)cudaErrorInvalidValue;
  __nv_bfloat16* d = reinterpret_cast<__nv_bfloat16*>(dst);
  const dim3 grid(tiles, l2norm ? 1 : (C + 63) / 64, N);
  if (l2norm && act) ws::nchw_to_keytiles_kernel<true, true><<<grid, 256, 0, st>>>(key, d, C, h, w, tw, th, tiles_x, slope, act_out);
  else if (l2norm) ws::nchw_to_keytiles_kernel<true, false><<<grid, 256, 0, st>>>(key, d, C, h, w, tw, th, tiles_x, 1.f, nullptr);
  else if (act) ws::nchw_to_keytiles_kernel<false, true><<<grid, 256, 0, st>>>(key, d, C, h, w, tw, th, tiles_x, slope, act_out);
  else ws::nchw_to_keytiles_kernel<false, false><<<grid, 256, 0, st>>>(key, d, C, h, w, tw, th, tiles_x, 1.f, nullptr);
  return (int)cudaGetLastError();
}

bool psweep_ws_eligible(const PsweepArgs& A, int nviews) {
  if (A.C % 32 != 0 || A.C > ws::kMaxC || A.ist[2] != 0 || A.ist[3] != 0) return false;
  if ((long long)A.h * A.w > (1 << 23)) return false;          // 32-bit element offsets inside a chunk of <= 256 planes
  for (int v = 0; v < nviews; ++v)
    if (A.hs[v] >= 32768 || A.ws[v] >= 32768) return false;
  return true;
}

int repack_grouped_batch(const float* const* srcs, void* const* dsts, const int* hs, const int* wss, int count, int N, int C, int l2norm,
                         unsigned* counter, cudaStream_t st, bool act, float slope) {
  if (count < 1 || count > kMaxViews || (long long)N * count > 65535) return (int)cudaErrorInvalidValue;
  ws::GroupBatch B;
  B.counter = counter;
  int gmax = 0;
  for (int i = 0; i < kMaxViews; ++i) {
    const int j = i < count ? i : 0;
    B.src[i] = srcs[j]; B.dst[i] = reinterpret_cast<__nv_bfloat16*>(dsts[j]); B.hs[i] = hs[j]; B.ws[i] = wss[j];
    if (i < count) {
      const int wg = (wss[i] + 7) >> 3, tiles = ((wg + 3) >> 2) * hs[i];
      if (tiles > gmax) gmax = tiles;
    }
  }
  const dim3 grid(gmax, l2norm ? 1 : (C + ws::kRpCh - 1) / ws::kRpCh, N * count);
  if (l2norm && act) ws::nchw_to_grouped_kernel<true, true><<<grid, 256, 0, st>>>(B, C, N, slope);
  else if (l2norm) ws::nchw_to_grouped_kernel<true, false><<<grid, 256, 0, st>>>(B, C, N, 1.f);
  else if (act) ws::nchw_to_grouped_kernel<false, true><<<grid, 256, 0, st>>>(B, C, N, slope);
  else ws::nchw_to_grouped_kernel<false, false><<<grid, 256, 0, st>>>(B, C, N, 1.f);
  return (int)cudaGetLastError();
}

int launch_psweep_ws(const PsweepArgs& A, const void* key_split, int nviews, unsigned* counter, const PsweepPeers* peers, cudaStream_t st) {
  static int num_sms = 0;
  static bool attr_set = false;
  if (!attr_set) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(ws::psweep_ws_kernel<ws::kModePlain>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ws::smem_bytes());
    if (e == cudaSuccess) e = cudaFuncSetAttribute(ws::psweep_ws_kernel<ws::kModePeers>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ws::smem_bytes());
    if (e == cudaSuccess) e = cudaFuncSetAttribute(ws::psweep_ws_kernel<ws::kModeDebug>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ws::smem_bytes());
    if (e == cudaSuccess) e = cudaFuncSetAttribute(ws::psweep_ws_kernel<ws::kModePush>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ws::smem_bytes());
    if (e != cudaSuccess) return (int)e;
    attr_set = true;
  }
  int tw, th;
  psweep_ws_tile_shape(A.h, A.w, &tw, &th);
  const int tiles_x = (A.w + tw - 1) / tw, tiles_y = (A.h + th - 1) / th;
  const int tiles = tiles_x * tiles_y;
  const long long items_ll = (long long)tiles * A.N * nviews;
  if (items_ll > 0x7fff0000LL) return (int)cudaErrorInvalidValue;
  int items = (int)items_ll;
  ws::PeerArgs PA{};
  PA.item_stride = 1;
  if (peers && peers->n > 0) {
    if (peers->n > ws::kMaxPeers || peers->rank < 0 || peers->rank >= peers->n || A.dbg_ix0) return (int)cudaErrorInvalidValue;
    PA.n = peers->n; PA.rank = peers->rank; PA.epoch = peers->epoch; PA.done_ctas = counter + 2;
    PA.corr_mc = peers->corr_mc; PA.mask_mc = peers->mask_mc;
    PA.q_head = peers->q_head; PA.q_cap = peers->q_cap;
    if (peers->by_tiles) {           // this rank's share of the items: rank, rank + n, rank + 2 n, ...
      PA.item_stride = peers->n; PA.item_offset = peers->rank;
      items = (items - peers->rank + peers->n - 1) / peers->n;
    }
    for (int g = 0; g < peers->n; ++g) { PA.corr[g] = peers->corr[g]; PA.mask[g] = peers->mask[g]; PA.flags[g] = peers->flags[g]; }
  }
  const bool push = PA.n > 0 && PA.q_head != nullptr;
  const int sms = push ? (num_sms - peers->push_ctas > 8 ? num_sms - peers->push_ctas : 8) : num_sms;   // push transport: the pusher kernel's SMs stay free
  const unsigned grid = (unsigned)(items < sms ? (items > 0 ? items : 1) : sms);   // a rank without items still runs the handshake
#ifdef RMVD_WS_STATS
  {
    unsigned long long z[32] = {0};
    cudaMemcpyToSymbolAsync(ws::g_ws_stats, z, sizeof(z), 0, cudaMemcpyHostToDevice, st);
  }
#endif
  const auto ks = reinterpret_cast<const __nv_bfloat16*>(key_split);
  if (push)
    ws::psweep_ws_kernel<ws::kModePush><<<grid, ws::kThreads, ws::smem_bytes(), st>>>(A, ks, tw, th, tiles_x, tiles, items, counter, PA);
  else if (PA.n > 0)
    ws::psweep_ws_kernel<ws::kModePeers><<<grid, ws::kThreads, ws::smem_bytes(), st>>>(A, ks, tw, th, tiles_x, tiles, items, counter, PA);
  else if (A.dbg_ix0)
    ws::psweep_ws_kernel<ws::kModeDebug><<<grid, ws::kThreads, ws::smem_bytes(), st>>>(A, ks, tw, th, tiles_x, tiles, items, counter, PA);
  else
    ws::psweep_ws_kernel<ws::kModePlain><<<grid, ws::kThreads, ws::smem_bytes(), st>>>(A, ks, tw, th, tiles_x, tiles, items, counter, PA);
#ifdef RMVD_WS_STATS
  {
    unsigned long long h[32];
    cudaStreamSynchronize(st);
    cudaMemcpyFromSymbol(h, ws::g_ws_stats, sizeof(h));
    const double c = (double)grid, ch = (double)(h[5] ? h[5] : 1);
    fprintf(stderr, "[psweep_ws] ctas=%u items=%d chunks=%llu (%.2f per item) mean_R=%.1f mean_len=%.1f band evals per chunk %.2f\n", grid, items, h[5],
            (double)h[5] / (double)items, (double)h[6] / ch, (double)h[7] / ch, (double)h[1] / ch);
    fprintf(stderr, "[psweep_ws] per CTA kcycles: planner total %.0f (wait cfree %.0f, band %.0f, wait plane table %.0f; table warp busy %.0f) | producer wait cready %.0f free %.0f | "
                    "mma wait cready %.0f aready %.0f drained %.0f full %.0f (stages %.0f)\n",
            h[4] / c / 1e3, h[0] / c / 1e3, h[2] / c / 1e3, h[3] / c / 1e3, h[15] / c / 1e3, h[8] / c / 1e3, h[9] / c / 1e3, h[10] / c / 1e3, h[11] / c / 1e3, h[12] / c / 1e3,
            h[13] / c / 1e3, h[14] / c);
    fprintf(stderr, "[psweep_ws] consumer (warp 4) per CTA kcycles: total %.0f = wait cready %.0f + bar1 %.0f + wait acc %.0f + drain %.0f + key %.0f + bar2 %.0f + "
                    "sample %.0f ; tiles %.1f\n",
            h[23] / c / 1e3, h[16] / c / 1e3, h[17] / c / 1e3, h[18] / c / 1e3, h[19] / c / 1e3, h[20] / c / 1e3, h[21] / c / 1e3, h[22] / c / 1e3, h[24] / c);
  }
#endif
  return (int)cudaGetLastError();
}

// q_shared[i] = a[i] / b[i] through rcp_refined() + div_shared() (the divisions of the exact coordinate chain), q_rn[i] = __fdiv_rn
__global__ void div_check_kernel(const float* __restrict__ a, const float* __restrict__ b, int n, float* __restrict__ q_shared,
                                 float* __restrict__ q_rn) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  q_shared[i] = div_shared(a[i], b[i], rcp_refined(b[i]));
  q_rn[i] = __fdiv_rn(a[i], b[i]);
}

}  // namespace rmvd

extern "C" int rmvd_debug_div_check(const float* a, const float* b, int n, float* q_shared, float* q_rn, void* stream) {
  if (!a || !b || !q_shared || !q_rn) return RMVD_E_NULL;
  if (n <= 0) return RMVD_E_SHAPE;
  rmvd::div_check_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(a, b, n, q_shared, q_rn);
  return (int)cudaGetLastError();
}
