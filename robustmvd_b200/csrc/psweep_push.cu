// Pusher of the depth-slab "push" transport (partition 1; no counterpart in the reference).
//
// The plane-sweep kernel (psweep_ws.cu, variant kModePush) stores its rows into this rank's own [V,N,S,h,w] tensors only and
// publishes one record per finished chunk of planes in a queue in device memory.  This kernel runs beside it on the few SMs the
// sweep leaves free (a CTA of 1024 threads cannot share an SM with a sweep CTA: registers), waits for the records (one per warp) and
// copies each chunk's rows -- `len` planes x th rows x tw pixels of corr and of mask -- from local memory (L2) to the same place
// in every peer's tensors with 16-byte loads and stores.  The remote stores stall THESE warps, not the sweep's consumer warps
// (peer stores from the sampling loop cost the sweep 0.24 ms at cfg4, 10 views, 2 GPUs: profiles/r02u_peer_local.jsonl).  The
// kernel ends with the flag handshake of the fused kernel: when it has completed on this rank's stream, every rank's slab has
// landed in this rank's tensors.
//
// Queue layout (unsigned words): [0] number of records of the call, [1] epoch of the call once [0] is final, [2] finished pusher
// CTAs, [3] unused; then q_cap flags (flag i = epoch once record i and its rows are visible); then q_cap records of 16 bytes
// {view-sample index, first output plane, planes, first pixel of the tile}.
#include "psweep.cuh"

namespace rmvd {

namespace {

struct PushArgs {
  const float* corr; const float* mask;
  float* pcorr[RMVD_MAX_PEERS]; float* pmask[RMVD_MAX_PEERS];
  unsigned* flags[RMVD_MAX_PEERS];
  unsigned* q_head; unsigned q_cap;
  int n, rank; unsigned epoch;
  int h, w, out_S, tw, th;
};

// A record is large (one chunk is often all planes of a tile: 128 planes x 64 pixels x 2 tensors = 64 KB at cfg4), so a CTA works on
// one record at a time, every warp on its share of the 16-byte units -- but the warps do not wait for each other: each follows
// the queue on its own (the flag and record reads hit L2) and looks one record ahead, so that the flag and record round trips
// of record i + gridDim.x are in flight while the rows of record i are copied.  CTA i takes records i, i + gridDim.x, ...
// (Measured before: one record per CTA with CTA-wide barriers 10 GB/s per SM, three dependent L2 round trips per record; one
// record per warp or 32 per warp: too few records -- 448 at cfg4 with 2 views -- to occupy the warps.  An SM's plain 16-byte
// peer stores carry 33 GB/s, 16 SMs 494 GB/s, 32 SMs 640 GB/s: profiles/r03i_peer_rate.jsonl.)
constexpr int kPushUnroll = 4;    // 16-byte units per lane in flight (corr + mask: 8 loads before the first store)
constexpr int kPushThreads = 512;   // (1024 threads leave 64 registers each: the units in flight spill)

__device__ __forceinline__ unsigned push_flag(const unsigned* qflags, unsigned idx, unsigned cap) {
  unsigned f = 0;
  if (idx < cap) asm volatile("ld.acquire.gpu.global.u32 %0, [%1];\n" : "=r"(f) : "l"(qflags + idx) : "memory");
  return f;
}

template <typename T>
__device__ __forceinline__ void push_record(const PushArgs& A, const int4& r, size_t P, int t) {
  constexpr int kPx = sizeof(T) / 4;
  const int x0 = r.w % A.w, y0 = r.w / A.w;
  const int width = min(A.tw, A.w - x0), rows = min(A.th, A.h - y0);
  const size_t base = ((size_t)r.x * A.out_S + r.y) * P + (size_t)y0 * A.w + x0;
  const int segs = width / kPx, per_plane = rows * segs, units = r.z * per_plane;
  for (int u0 = t; u0 < units; u0 += kPushThreads * kPushUnroll) {
    T c[kPushUnroll], m[kPushUnroll];
    size_t o[kPushUnroll];
#pragma unroll
    for (int k = 0; k < kPushUnroll; ++k) {
      const int u = u0 + kPushThreads * k;
      if (u < units) {
        const int pl = u / per_plane, q = u - pl * per_plane, ty = q / segs, sg = q - ty * segs;
        o[k] = base + (size_t)pl * P + (size_t)ty * A.w + kPx * sg;
        c[k] = __ldcg(reinterpret_cast<const T*>(A.corr + o[k]));
        m[k] = __ldcg(reinterpret_cast<const T*>(A.mask + o[k]));
      }
    }
    for (int p = 0; p < A.n; ++p) {
      if (p == A.rank) continue;
      float* const pc = A.pcorr[p];
      float* const pm = A.pmask[p];
#pragma unroll
      for (int k = 0; k < kPushUnroll; ++k) {
        if (u0 + kPushThreads * k < units) { *reinterpret_cast<T*>(pc + o[k]) = c[k]; *reinterpret_cast<T*>(pm + o[k]) = m[k]; }
      }
    }
  }
}

__global__ void __launch_bounds__(kPushThreads, 1) psweep_push_kernel(const __grid_constant__ PushArgs A) {
  const int t = threadIdx.x;
  const size_t P = (size_t)A.h * A.w;
  const unsigned* qflags = A.q_head + 4;
  const int4* recs = reinterpret_cast<const int4*>(A.q_head + 4 + A.q_cap);
  const bool vec = ((A.w | A.tw) & 3) == 0;            // tiles start at multiples of tw: every row segment is 16-byte aligned
  const unsigned stride = gridDim.x;
  unsigned i = blockIdx.x;
  // every lane reads the same words (one broadcast transaction per warp): no shuffles, and each lane's own acquire orders its reads
  unsigned f_cur = push_flag(qflags, i, A.q_cap), f_nxt = push_flag(qflags, i + stride, A.q_cap);
  int4 rec = make_int4(0, 0, 0, 0);
  bool have_rec = false;
  for (;;) {
    if (f_cur != A.epoch) {                              // not published when we looked ahead: wait for it, or for the end of the call
      bool done = false;
      unsigned polls = 0;
      for (;; ++polls) {
        f_cur = push_flag(qflags, i, A.q_cap);
        if (f_cur == A.epoch) break;
        unsigned fin, total;
        asm volatile("ld.acquire.gpu.global.u32 %0, [%1];\n" : "=r"(fin) : "l"(A.q_head + 1) : "memory");
        asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];\n" : "=r"(total) : "l"(A.q_head) : "memory");
        if (fin == A.epoch && i >= total) { done = true; break; }
        if (polls > (1u << 24)) __trap();                // several seconds: a sweep that never publishes ends the wait
        __nanosleep(100);
      }
      if (done) break;
      have_rec = false;
    }
    if (!have_rec) rec = __ldcg(recs + i);
    const bool nxt = f_nxt == A.epoch;                   // look ahead: the next record's fields and the flag after it
    int4 rec_n = make_int4(0, 0, 0, 0);
    if (nxt) rec_n = __ldcg(recs + i + stride);
    const unsigned f_nn = push_flag(qflags, i + 2 * stride, A.q_cap);
    if (vec) push_record<float4>(A, rec, P, t);
    else push_record<float>(A, rec, P, t);
    i += stride;
    f_cur = f_nxt; f_nxt = f_nn; rec = rec_n; have_rec = nxt;
  }
  __threadfence_system();                                // this thread's peer stores are ordered before the handshake
  __syncthreads();
  if (t == 0) {
    const unsigned prev = atomicAdd(A.q_head + 2, 1u);
    if (prev == gridDim.x - 1) {
      A.q_head[2] = 0u;
      __threadfence_system();
      for (int g = 0; g < A.n; ++g) asm volatile("st.release.sys.global.u32 [%0], %1;\n" ::"l"(A.flags[g] + A.rank), "r"(A.epoch) : "memory");
      for (int g = 0; g < A.n; ++g) {
        unsigned have = 0, polls = 0;
        for (;;) {
          asm volatile("ld.acquire.sys.global.u32 %0, [%1];\n" : "=r"(have) : "l"(A.flags[A.rank] + g) : "memory");
          if ((int)(have - A.epoch) >= 0) break;
          __nanosleep(200);
          if (++polls > (1u << 24)) __trap();         // a peer never arrived (several seconds): fail loudly rather than hang
        }
      }
    }
  }
}

}  // namespace

int launch_psweep_push(const PsweepPeers& peers, int h, int w, int out_S, cudaStream_t st) {
  if (!peers.q_head || peers.n < 1 || peers.n > RMVD_MAX_PEERS || peers.push_ctas < 1) return (int)cudaErrorInvalidValue;
  PushArgs A{};
  A.corr = peers.corr[peers.rank]; A.mask = peers.mask[peers.rank];
  for (int g = 0; g < peers.n; ++g) { A.pcorr[g] = peers.corr[g]; A.pmask[g] = peers.mask[g]; A.flags[g] = peers.flags[g]; }
  A.q_head = peers.q_head; A.q_cap = peers.q_cap; A.n = peers.n; A.rank = peers.rank; A.epoch = peers.epoch;
  A.h = h; A.w = w; A.out_S = out_S;
  psweep_ws_tile_shape(h, w, &A.tw, &A.th);
  psweep_push_kernel<<<peers.push_ctas, kPushThreads, 0, st>>>(A);
  return (int)cudaGetLastError();
}

}  // namespace rmvd
