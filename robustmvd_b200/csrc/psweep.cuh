// Argument block shared by the plane-sweep correlation kernels (psweep.cu: general direct kernel,
// psweep_band.cu: banded fp32 kernel, psweep_ws.cu: persistent tcgen05 kernel).
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
  int border;                        // kFromGrid: grid_sample padding_mode="border" (coordinates clipped into the map, mask = geometric mask only)
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

// psweep_ws.cu: persistent warp-specialised tcgen05 kernel (one CTA per SM).  The key map is repacked tile by tile
// (repack_key_tiles: psweep_ws_key_bytes), the source maps into the grouped split layout of repack_grouped_batch() (rows
// padded to multiples of 8 texels: psweep_ws_src_bytes).
bool psweep_ws_eligible(const PsweepArgs& A, int nviews);
size_t psweep_ws_src_bytes(int N, int C, int hs, int ws);
size_t psweep_ws_key_bytes(int N, int C, int h, int w);
int repack_key_tiles(const float* key, void* dst, int N, int C, int h, int w, int l2norm, cudaStream_t st, bool act = false, float slope = 1.f,
                     float* act_out = nullptr);
// `counter`: 4 bytes of device memory (the kernel's work-item counter), zeroed by the repack launch
int repack_grouped_batch(const float* const* srcs, void* const* dsts, const int* hs, const int* ws, int count, int N, int C, int l2norm,
                         unsigned* counter, cudaStream_t st, bool act = false, float slope = 1.f);
// Depth-slab mode: the sampling loop stores every output row into the tensors of all ranks (peer-mapped pointers, own rank
// included) and the kernel ends with a flag handshake between the ranks.
struct PsweepPeers {
  int n, rank;
  float* corr[RMVD_MAX_PEERS]; float* mask[RMVD_MAX_PEERS];
  unsigned* flags[RMVD_MAX_PEERS];
  unsigned epoch;
  float* corr_mc; float* mask_mc;  // optional NVSwitch multicast addresses of the same tensors (multimem.st: one store reaches every rank); NULL: one store per rank
  unsigned* q_head; unsigned q_cap; int push_ctas;   // push transport: chunk queue (NULL: stores from the sampling loop), SMs left to the pusher kernel
  int by_tiles;                    // 0: the ranks split the planes (the caller's [s_begin, s_end)); 1: they split the work items (tiles), all planes each
};
// counter: 16 bytes of device memory: [0] work-item counter, [2] finished-CTA counter of the slab handshake
int launch_psweep_ws(const PsweepArgs& A, const void* key_split, int nviews, unsigned* counter, const PsweepPeers* peers, cudaStream_t st);
// Pusher kernel of the depth-slab push transport: forwards the chunks the plane-sweep kernel publishes in the queue to every peer and
// ends with the ranks' flag handshake.  corr / mask: this rank's own [V,N,out_S,h,w] tensors (entry `rank` of peers->corr / mask).
int launch_psweep_push(const PsweepPeers& peers, int h, int w, int out_S, cudaStream_t st);
void psweep_ws_tile_shape(int h, int w, int* tw_out, int* th_out);   // pixel tile of psweep_ws_kernel for an h x w key image

}  // namespace rmvd
