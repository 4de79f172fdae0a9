/*
 * rmvd_b200.h -- C ABI of the B200-native plane-sweep cost-volume path.
 *
 * The reference (saiprasadbarke/robustmvd) has no FFI layer: its boundary for this path is the Python
 * call signature of the blocks.  Every entry point below replaces one of those Python call sites (cited as
 * reference file:line) and is what a reference-side binding (ctypes stub) would bind.
 *
 * Conventions
 *   - all tensors float32, contiguous, NCHW unless stated; pointers are DEVICE pointers unless the name
 *     ends in `_host` or the comment says "host array";
 *   - every call is asynchronous on `stream` (a cudaStream_t passed as void*), re-entrant, and holds no
 *     global mutable state; scratch memory is passed in by the caller (`workspace`) -- the exceptions say so:
 *     rmvd_psweep_corr_fwd_static (static device scratch) and rmvd_host_expand_mask[_async] (host worker pool);
 *   - return value: 0 ok, <0 invalid argument (RMVD_E_*), >0 a cudaError_t; nothing throws across the ABI;
 *   - there is no CPU fallback: without a CUDA device the calls return a cudaError_t.
 */
#ifndef RMVD_B200_H_
#define RMVD_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RMVD_ABI_VERSION 8
#define RMVD_MAX_PEERS 8      /* ranks of one depth-slab communicator (GPUs of one NVSwitch box) */

#define RMVD_OK 0
#define RMVD_E_NULL (-1)      /* required pointer is NULL */
#define RMVD_E_SHAPE (-2)     /* non-positive or inconsistent size */
#define RMVD_E_MODE (-3)      /* unknown enum value */
#define RMVD_E_WORKSPACE (-4) /* workspace too small */
#define RMVD_E_RANGE (-5)     /* plane range outside [0, S] */
#define RMVD_E_STATIC_LIMIT (-6) /* problem needs more scratch than RMVD_STATIC_SCRATCH_BYTES (rmvd_psweep_corr_fwd_static) */

/* normalize= of TorchCorr / WarpOnlyCorr (planesweep_corr.py:108-114, :144-150) */
#define RMVD_NORM_NONE 0   /* False */
#define RMVD_NORM_DIM 1    /* "dim": corr / sqrt(C) (TorchCorr); no-op for WarpOnlyCorr */
#define RMVD_NORM_BEFORE 2 /* True or "before": L2-normalise features along C first */
#define RMVD_NORM_AFTER 3  /* WarpOnlyCorr only ("after" or True): L2-normalise the warped features */

int rmvd_abi_version(void);
/* Human-readable name of a status returned by the calls below (static storage). */
const char* rmvd_status_string(int status);

/* ---------------------------------------------------------------------------------------------------
 * robust_mvd family
 * ------------------------------------------------------------------------------------------------- */

/* Scratch bytes rmvd_psweep_corr_fwd needs (channels-last copies of the key and the V source feature maps). */
size_t rmvd_psweep_workspace_bytes(int N, int C, int h, int w, int V, const int* hs_host, const int* ws_host);

/*
 * PlanesweepCorrelation.forward with warp_only=False (planesweep_corr.py:396-427 -> :436-521) fused into one
 * pass per source view: EpipolarCoeffs.from_calib (:228-300), us_from_ds/vs_from_ds (:333-349),
 * get_plane_sweep_sampling_points (:489-512), TorchCorr.forward (:152-195) and warp's grid_sample + sampling
 * mask (:85-102).  The all-pairs matrix of :174 is never formed: corr = mask/sqrt(C) * <key, bilinear(src)>.
 *
 *   feat_key        [N,C,h,w]
 *   feat_src        host array of V device pointers, view v is [N,C,hs[v],ws[v]]
 *   intrinsics_key  [N,3,3] RELATIVE intrinsics (fx/W ...), robust_mvd.py:117-120
 *   intrinsics_src  host array of V device pointers to [N,3,3]
 *   src_to_key      host array of V device pointers to [N,4,4] (p_src = T p_key)
 *   invdepths       sampling inverse depths; element (n,s,y,x) is read at
 *                   invdepths[n*inv_stride[0] + s*inv_stride[1] + y*inv_stride[2] + x*inv_stride[3]]
 *                   (strides in elements, 0 for broadcast dims: covers (1|N, S), (N,S,H), (N,S,H,W), :483-485)
 *   planes [s_begin, s_end) of the S hypotheses are computed (depth-slab partition); plane s is written at
 *   index out_s0 + (s - s_begin) of outputs shaped [V,N,out_S,h,w].
 *   corr, mask      float32 [V,N,out_S,h,w]; mask holds 0/1 like the reference's corr_mask (:189-193)
 *   dbg_ix0,dbg_iy0 optional int32 [V,N,out_S,h,w]: floor of the unnormalised sampling coordinate,
 *                   saturated to [-2, ws] / [-2, hs]; dbg_geom optional uint8: geometric mask (:499-506)
 */
int rmvd_psweep_corr_fwd(const float* feat_key, const float* const* feat_src, const float* intrinsics_key,
                         const float* const* intrinsics_src, const float* const* src_to_key,
                         const float* invdepths, const int64_t* inv_stride_host, int N, int C, int h, int w,
                         int V, const int* hs_host, const int* ws_host, int S, int s_begin, int s_end,
                         int normalize_mode, float* corr, float* mask, int out_S, int out_s0,
                         int32_t* dbg_ix0, int32_t* dbg_iy0, uint8_t* dbg_geom, void* workspace,
                         size_t workspace_bytes, void* stream);

/*
 * Depth-slab form of rmvd_psweep_corr_fwd for G = num_peers GPUs of one NVLink / NVSwitch box (partition 1; the
 * reference has no counterpart -- its only multi-GPU mode is nn.DataParallel over the batch, rmvd/models/helpers.py:161-170).
 * Rank `rank` computes planes [s_begin, s_end) and its sampling loop stores every corr / mask row into the [V,N,out_S,h,w]
 * tensors of ALL ranks (corr_peers[g], mask_peers[g]: peer-mapped device pointers, entry `rank` = this rank's own tensors), at
 * plane index out_s0 + (s - s_begin): the all-gather that the consumer (LearnedFusion, learned_fusion.py:13-17: planes are
 * channels) needs happens inside the kernel over NVLink.  The kernel ends with a handshake: the last CTA writes `epoch` to slot
 * `rank` of every rank's flag array (flag_peers[g]: peer-mapped uint32[num_peers], zero before the first call) and waits until
 * its own array holds >= epoch in every slot.  Hence, when the call's kernel has completed on this rank's stream, all ranks'
 * slabs are present in this rank's tensors.  `epoch` must increase by one per call of the communicator; callers alternate
 * between two sets of output tensors so that a rank that runs ahead never overwrites rows a slower rank is still reading.
 * All ranks must issue the call (same shapes, own slab); num_peers <= RMVD_MAX_PEERS; the debug outputs are not available.
 * Requires the persistent tcgen05 kernel (C % 32 == 0, C <= 256, per-plane hypotheses): RMVD_E_MODE otherwise.
 * partition: RMVD_PARTITION_PLANES -- rank computes the planes [s_begin, s_end) of every tile (the depth slabs of SURVEY.md 8e);
 *            RMVD_PARTITION_TILES  -- every rank passes the same plane range (normally all S planes) and computes every
 *            num_peers-th (tile, sample, view) work item of it: the same tensors result on every rank, but the per-tile costs that
 *            do not depend on the plane count (key tile, calibration, band planning, the band's rows) are divided between the
 *            ranks too, which the plane split cannot do.
 * corr_multicast / mask_multicast (both or neither): NVSwitch multicast addresses (NVLS) of the same symmetric tensors; the sampling
 *            loop then issues ONE multimem.st per value, which the switch replicates into every rank's tensor (own rank included),
 *            instead of num_peers stores: the SM's egress falls from (num_peers - 1) to 1 copy of its slab.
 */
#define RMVD_PARTITION_PLANES 0
#define RMVD_PARTITION_TILES 1
int rmvd_psweep_corr_fwd_peers(const float* feat_key, const float* const* feat_src, const float* intrinsics_key,
                               const float* const* intrinsics_src, const float* const* src_to_key,
                               const float* invdepths, const int64_t* inv_stride_host, int N, int C, int h, int w,
                               int V, const int* hs_host, const int* ws_host, int S, int s_begin, int s_end,
                               int normalize_mode, int out_S, int out_s0, int num_peers, int rank,
                               float* const* corr_peers, float* const* mask_peers, uint32_t* const* flag_peers,
                               uint32_t epoch, int partition, float* corr_multicast, float* mask_multicast,
                               void* workspace, size_t workspace_bytes, void* stream);

/*
 * Depth slabs with a PUSHER kernel as the transport (no counterpart in the reference, see rmvd_psweep_corr_fwd_peers).  The sweep
 * (on `stream`) stores its planes [s_begin, s_end) into this rank's own tensors corr_peers[rank] / mask_peers[rank] and publishes
 * every finished chunk in `queue`; a second kernel (on `push_stream`, push_ctas CTAs on SMs the sweep leaves free) forwards the
 * chunks' rows to the same place in every peer's tensors with 16-byte stores and ends with the ranks' flag handshake: when the work
 * of `push_stream` has completed, all ranks' slabs are present in this rank's tensors.  The remote stores stall the pusher's warps
 * instead of the sweep's sampling warps.  queue: 16-byte aligned device memory of rmvd_psweep_push_queue_bytes(N, h, w, V, planes of
 * the slab) bytes, zeroed once when it is allocated, owned by the communicator (one call at a time); `epoch` as in
 * rmvd_psweep_corr_fwd_peers.  stream != push_stream; the caller makes `stream` wait for `push_stream` before the next call.
 */
size_t rmvd_psweep_push_queue_bytes(int N, int h, int w, int V, int planes);
int rmvd_psweep_corr_fwd_push(const float* feat_key, const float* const* feat_src, const float* intrinsics_key,
                              const float* const* intrinsics_src, const float* const* src_to_key,
                              const float* invdepths, const int64_t* inv_stride_host, int N, int C, int h, int w,
                              int V, const int* hs_host, const int* ws_host, int S, int s_begin, int s_end,
                              int normalize_mode, int out_S, int out_s0, int num_peers, int rank,
                              float* const* corr_peers, float* const* mask_peers, uint32_t* const* flag_peers,
                              uint32_t epoch, void* queue, size_t queue_bytes, int push_ctas, void* workspace,
                              size_t workspace_bytes, void* stream, void* push_stream);

/*
 * Transport of the depth-slab gather by the copy engines (no counterpart in the reference, see rmvd_psweep_corr_fwd_peers): `rows`
 * runs of row_bytes bytes, pitch_bytes apart in both buffers, from this rank's tensor to the same place in a peer-mapped tensor
 * (one cudaMemcpy2DAsync).  Used by parallel.SlabGather(transport="copy"): the slab of view group k travels while the kernel of
 * group k + 1 runs.
 */
int rmvd_copy_rows(void* dst, const void* src, size_t pitch_bytes, size_t row_bytes, int rows, void* stream);

/*
 * rmvd_psweep_corr_fwd without any allocation by anyone: the scratch (what rmvd_psweep_workspace_bytes() sizes) is a
 * file-scope device array of RMVD_STATIC_SCRATCH_BYTES inside the library instead of a caller's workspace.  This entry
 * point has no counterpart in the reference; it was added for callers that cannot
 * call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The call itself
 * issues no cudaMalloc / cudaFree and no allocating library routine; the caller passes inputs and outputs as for
 * rmvd_psweep_corr_fwd (same meaning, same results; no debug outputs).
 *   limit     rmvd_psweep_workspace_bytes(N, C, h, w, V, hs, ws) <= RMVD_STATIC_SCRATCH_BYTES (32 MiB: one frame of the
 *             KITTI shape, C = 256, 48 x 160 features, with up to 3 source views = 31.5 MB); larger problems return
 *             RMVD_E_STATIC_LIMIT and launch nothing -- use rmvd_psweep_corr_fwd with a workspace.
 *   cost      every GPU that loads the library holds the 32 MiB until the process exits.
 *   sharing   all calls of a process on one GPU use the same array, which keeps the previous call's contents (each call
 *             zeroes the counters it needs and overwrites the rest before reading it): calls that overlap on the GPU
 *             -- different streams, or different threads -- race.  Serialise them on one stream.
 *   device    the array's device address is looked up once per device (cudaGetSymbolAddress with the calling thread's
 *             current device) and cached; call with the device of the tensors current.
 */
#define RMVD_STATIC_SCRATCH_BYTES (32u << 20)
int rmvd_psweep_corr_fwd_static(const float* feat_key, const float* const* feat_src, const float* intrinsics_key,
                                const float* const* intrinsics_src, const float* const* src_to_key,
                                const float* invdepths, const int64_t* inv_stride_host, int N, int C, int h, int w,
                                int V, const int* hs_host, const int* ws_host, int S, int s_begin, int s_end,
                                int normalize_mode, float* corr, float* mask, int out_S, int out_s0, void* stream);

/*
 * rmvd_psweep_corr_fwd with the activation of the feature encoder's last layer fused into the emission of the kernel's operand
 * layouts (second half).  The reference computes the features with Conv2d + LeakyReLU(0.2)
 * (dispnet_encoder.py:23-27 through blocks/utils.py:14-28; robust_mvd.py:66-67) and hands the activated NCHW maps to the block; here
 * the caller passes the convolution's PRE-activation outputs and the split copies (bf16 hi / lo, tile-ordered key map, grouped
 * source maps) apply x > 0 ? x : x * act_slope as they read them: the activated float32 source maps are never materialised and
 * the separate activation pass disappears.
 *   pre_key, pre_src   as feat_key / feat_src of rmvd_psweep_corr_fwd, before the activation
 *   act_slope          negative slope (0.2 for the DispNet encoder; 1.0 = identity, 0.0 = ReLU)
 *   key_act_out        [N,C,h,w] or NULL: receives the activated key map (it has other consumers: dispnet_context_encoder.py:12,
 *                      the decoder's skip connection); must not overlap pre_key
 * Results are bit-identical to rmvd_psweep_corr_fwd on the activated maps.  Needs the persistent kernel's shapes (C % 32 == 0,
 * C <= 256, per-plane hypotheses); otherwise RMVD_E_MODE and the caller activates first.  No debug outputs.
 */
int rmvd_psweep_corr_fwd_act(const float* pre_key, const float* const* pre_src, float act_slope, float* key_act_out,
                             const float* intrinsics_key, const float* const* intrinsics_src, const float* const* src_to_key,
                             const float* invdepths, const int64_t* inv_stride_host, int N, int C, int h, int w,
                             int V, const int* hs_host, const int* ws_host, int S, int s_begin, int s_end,
                             int normalize_mode, float* corr, float* mask, int out_S, int out_s0, void* workspace,
                             size_t workspace_bytes, void* stream);

/*
 * Compact transport of the sampling mask for callers whose outputs go back to HOST memory (pipeline.HostPlanesweepPipeline; the
 * reference keeps corr and mask on the device, robust_mvd.py:71-82, so these three have no counterpart there).  The mask of
 * PlanesweepCorrelation.forward (planesweep_corr.py:99-102, :189-191) is a float32 tensor of 0.0 / 1.0 values: half of the bytes
 * of a device-to-host copy of the block's outputs.
 *   rmvd_mask_pack_u8            device: packed[i] = mask[i] != 0 (one byte per element; mask 16-byte, packed 4-byte aligned)
 *   rmvd_host_expand_mask        host, synchronous: mask_host[i] = (float)packed_host[i] on `threads` worker threads (streaming
 *                                stores); the result is bit-identical to a float32 copy of the device mask
 *   rmvd_host_expand_mask_async  the same as a host function queued on `stream` (cudaLaunchHostFunc): runs after the work queued
 *                                before it (the device-to-host copy of `packed`), work queued after it on the stream waits for it
 * Both host pointers must stay valid until the expansion has run.  The worker pool (<= 32 threads, created on first use) is the one
 * piece of process-wide state; concurrent expansions are serialised.
 */
int rmvd_mask_pack_u8(const float* mask, uint8_t* packed, size_t n, void* stream);
int rmvd_host_expand_mask(const uint8_t* packed_host, float* mask_host, size_t n, int threads);
int rmvd_host_expand_mask_async(const uint8_t* packed_host, float* mask_host, size_t n, int threads, void* stream);
/* The same three with ONE BIT per element (word j, bit i = element 32 j + i; ceil(n / 32) words; little-endian host): an eighth of
 * the bytes again -- what pipeline.HostPlanesweepPipeline uses by default. */
int rmvd_mask_pack_bits(const float* mask, uint32_t* packed, size_t n, void* stream);
int rmvd_host_expand_mask_bits(const uint32_t* packed_host, float* mask_host, size_t n, int threads);
int rmvd_host_expand_mask_bits_async(const uint32_t* packed_host, float* mask_host, size_t n, int threads, void* stream);

/*
 * Backward of rmvd_psweep_corr_fwd with respect to the feature maps.  The reference obtains it from
 * autograd through TorchCorr.forward (planesweep_corr.py:152-195: matmul -> / sqrt(C) -> grid_sample -> * corr_mask ->
 * * geometric mask); sampling points carry no gradient (:436, :464, :489 are no_grad).  Same geometry arguments as the
 * forward call; normalize_mode RMVD_NORM_NONE or RMVD_NORM_DIM (L2-normalised inputs are differentiated by the caller
 * around a RMVD_NORM_NONE call).
 *   grad_corr  [V,N,out_S,h,w], plane s of [s_begin, s_end) read at index out_s0 + (s - s_begin)
 *   grad_key   [N,C,h,w] (overwritten) or NULL
 *   grad_src   host array of V device pointers, view v is [N,C,hs[v],ws[v]] (overwritten); entries may be NULL
 *   workspace  rmvd_psweep_bwd_workspace_bytes(...) bytes of device scratch
 */
size_t rmvd_psweep_bwd_workspace_bytes(int N, int C, int h, int w, int V, const int* hs_host, const int* ws_host);
int rmvd_psweep_corr_bwd(const float* feat_key, const float* const* feat_src, const float* intrinsics_key,
                         const float* const* intrinsics_src, const float* const* src_to_key,
                         const float* invdepths, const int64_t* inv_stride_host, int N, int C, int h, int w,
                         int V, const int* hs_host, const int* ws_host, int S, int s_begin, int s_end,
                         int normalize_mode, const float* grad_corr, int out_S, int out_s0, float* grad_key,
                         float* const* grad_src, void* workspace, size_t workspace_bytes, void* stream);

/*
 * PlanesweepCorrelation.forward with warp_only=True: WarpOnlyCorr.forward + warp_multi
 * (planesweep_corr.py:107-140, :13-46).  One source view per call.
 *   warped [N,out_S,C,h,w], mask [N,out_S,h,w] (sampling mask only: the geometric mask is ignored, :116-140)
 */
int rmvd_psweep_warp_fwd(const float* feat_src, const float* intrinsics_key, const float* intrinsics_src,
                         const float* src_to_key, const float* invdepths, const int64_t* inv_stride_host,
                         int N, int C, int h, int w, int hs, int ws, int S, int s_begin, int s_end,
                         int normalize_mode, float* warped, float* mask, int out_S, int out_s0, void* stream);

/*
 * Backward of rmvd_psweep_warp_fwd with respect to the source feature map (training direction of WarpOnlyCorr; the reference gets it
 * from autograd through grid_sample and `* warping_mask`, planesweep_corr.py:131-138).  normalize_mode RMVD_NORM_NONE or
 * RMVD_NORM_DIM (the L2-normalised variants are differentiated by the caller around a RMVD_NORM_NONE call).  Same geometry arguments
 * as the forward call; grad_warped [N,out_S,C,h,w] (plane s read at out_s0 + s - s_begin), grad_src [N,C,hs,ws] (zeroed by the call,
 * float atomics).  Sampling positions carry no gradient.
 */
int rmvd_psweep_warp_bwd(const float* intrinsics_key, const float* intrinsics_src, const float* src_to_key,
                         const float* invdepths, const int64_t* inv_stride_host, int N, int C, int h, int w, int hs, int ws,
                         int S, int s_begin, int s_end, int normalize_mode, const float* grad_warped, int out_S, int out_s0,
                         float* grad_src, void* stream);

/*
 * The state the reference block keeps after forward(): sampling points of one source view (planesweep_corr.py:489-512) --
 * us, vs (source pixel coordinates, non-finite values replaced as :336-338), zs = 1 / invdepth, geom_mask (uint8, :499-506), all
 * [N,S,h,w] -- and its epipolar coefficients (:228-300): coeff_pix [N,3,h,w] = (u_infs_h, v_infs_h, k_infs_h), coeff_m [N,3] =
 * (m_u_h, m_v_h, m_k_h).  Either group may be NULL.  The fused correlation never materialises these; the Python block computes
 * them lazily when `block.sampling_points[i].us` / `block.coeffs[i]` is read.
 */
int rmvd_sampling_points_fwd(const float* intrinsics_key, const float* intrinsics_src, const float* src_to_key,
                             const float* invdepths, const int64_t* inv_stride_host, int N, int h, int w, int hs, int ws,
                             int S, float* us, float* vs, float* zs, uint8_t* geom_mask, float* coeff_pix, float* coeff_m,
                             void* stream);

/*
 * TorchCorr.forward on caller-supplied sampling grids (planesweep_corr.py:152-195): grids [N,S,2,h,w] in
 * source pixels, geom_mask optional uint8 [N,S,h,w].  Outputs corr, mask float32 [N,S,h,w].
 * padding_border: TorchCorr's constructor argument (:144-148): 0 = "zeros" (the block's use), 1 = "border" (coordinates clipped
 * into the source map as ATen does, sampling mask all ones: mask = geom_mask).
 */
int rmvd_corr_grid_fwd(const float* feat_key, const float* feat_src, const float* grids,
                       const uint8_t* geom_mask, int N, int C, int h, int w, int hs, int ws, int S,
                       int normalize_mode, int padding_border, float* corr, float* mask, void* workspace,
                       size_t workspace_bytes, void* stream);

/*
 * warp() grid branch (planesweep_corr.py:49-104): x [K / x_repeat,C,hs,ws], grid [K,2,ho,wo] in source pixels,
 * padding_border: 0 = "zeros", 1 = "border".  out [K,C,ho,wo], mask [K,1,ho,wo].  x_repeat = 1: warp(); x_repeat = S with
 * K = N * S: warp_multi() (:13-46) in one launch, grids k = n * S + s all sample image n (the reference's repeat of x, :33, is
 * not materialised).
 */
int rmvd_warp_grid_fwd(const float* x, const float* grid, int K, int C, int hs, int ws, int ho, int wo,
                       int padding_border, int x_repeat, float* out, float* mask, void* stream);

/* ---------------------------------------------------------------------------------------------------
 * MVSNet family
 * ------------------------------------------------------------------------------------------------- */

/*
 * MVSNet cost-volume assembly (mvsnet.py:123-136) with homo_warp (blocks/utils.py:222-268) fused in: the
 * warped source volumes are never written.  feats [B,Vt,C,h,w] (view 0 = key), proj [B,Vt,4,4] (view 0
 * already inverted, mvsnet.py:85-86), depths [B,D].  Planes [d_begin,d_end) -> variance [B,C,out_D,h,w] at
 * plane index out_d0 + (d - d_begin).
 */
int rmvd_mvs_variance_fwd(const float* feats, const float* proj, const float* depths, int B, int Vt, int C,
                          int h, int w, int D, int d_begin, int d_end, float* variance, int out_D, int out_d0,
                          void* stream);

/*
 * The same volume (mvsnet.py:123-136 + blocks/utils.py:222-268) through the channels-last staged kernel (csrc/mvs_var.cu):
 * the source views are first repacked into `workspace` ([B,Vt-1,h*w,C] float32), a staged texel is one 128-byte line and a
 * corner 8 x LDS.128.  Takes C % 32 == 0 and Vt >= 2; every other shape is forwarded to rmvd_mvs_variance_fwd (no
 * workspace needed: rmvd_mvs_variance_workspace_bytes returns 0 for it).  workspace: 16-byte aligned device memory of at
 * least rmvd_mvs_variance_workspace_bytes(B, Vt, C, h, w) bytes, scratch (contents undefined afterwards).
 */
size_t rmvd_mvs_variance_workspace_bytes(int B, int Vt, int C, int h, int w);
int rmvd_mvs_variance_fwd_ws(const float* feats, const float* proj, const float* depths, int B, int Vt, int C, int h,
                             int w, int D, int d_begin, int d_end, float* variance, int out_D, int out_d0,
                             void* workspace, size_t workspace_bytes, void* stream);

/*
 * homo_warp (blocks/utils.py:222-268), materialising: src_feat [B,C,H,W], src_proj/ref_proj_inv [B,4,4],
 * depths [B,D] -> warped [B,C,D,H,W]; dbg_ix0/dbg_iy0 optional int32 [B,D,H,W] (saturated to [-2,W]/[-2,H]).
 */
int rmvd_homo_warp_fwd(const float* src_feat, const float* src_proj, const float* ref_proj_inv,
                       const float* depths, int B, int C, int H, int W, int D, float* warped,
                       int32_t* dbg_ix0, int32_t* dbg_iy0, void* stream);

/*
 * softmax over D + depth_regression + 4-bin confidence (mvsnet.py:140-160, blocks/utils.py:271-274) in one
 * pass.  cost [B,D,h,w], depths [B,D].  Outputs (each optional) depth, confidence float32 [B,h,w],
 * depth_index int64 [B,h,w] (= trunc(sum_d p_d * d), clamped to [0,D-1]), expected_index float32 [B,h,w],
 * prob float32 [B,D,h,w].
 */
int rmvd_mvs_regress_fwd(const float* cost, const float* depths, int B, int D, int h, int w, float* depth,
                         float* confidence, int64_t* depth_index, float* expected_index, float* prob,
                         void* stream);

/*
 * Backward of the MVSNet cost-volume path (training direction; extends SURVEY.md row 8f-4 to the MVSNet family).  The reference
 * obtains both from autograd: through homo_warp's grid_sample (blocks/utils.py:262) and the variance expression
 * (mvsnet.py:123-136), and through F.softmax + depth_regression (mvsnet.py:140-141, blocks/utils.py:271-274).  Sampling positions
 * carry no gradient (depth hypotheses and projections are data); the confidence is computed under no_grad (mvsnet.py:143-158).
 *   rmvd_mvs_variance_bwd  grad_variance [B,C,D,h,w] -> grad_feats [B,Vt,C,h,w] (zeroed by the call; view 0 = key view).
 *                          d var / d x_v = (2 / Vt) (x_v - mean) with x_0 = ref, x_v = warp_v(src_v): the warps are recomputed,
 *                          source views receive the transpose of their bilinear gather (float atomics: sums in any order)
 *   rmvd_mvs_regress_bwd   grad_depth [B,h,w] -> grad_cost [B,D,h,w]: g * softmax(cost)_d * (depth_d - depth)
 */
int rmvd_mvs_variance_bwd(const float* feats, const float* proj_mats, const float* depth_values, const float* grad_variance,
                          int B, int Vt, int C, int h, int w, int D, float* grad_feats, void* stream);
int rmvd_mvs_regress_bwd(const float* cost, const float* depth_values, const float* grad_depth, int B, int D, int h, int w,
                         float* grad_cost, void* stream);

/*
 * Depth-slab softargmax.  Step 1 on each rank: partials over planes [d_begin,d_end) of a
 * slab cost [B,Dl,h,w] (Dl = d_end - d_begin), depths_slab [B,Dl]:
 *   partial [4,B,h,w] = (m = max, Z = sum exp(x-m), A = sum exp(x-m)*depth, I = sum exp(x-m)*index).
 */
int rmvd_softargmax_partial(const float* cost_slab, const float* depths_slab, int B, int Dl, int h, int w,
                            int d_begin, float* partial, void* stream);
/*
 * Step 2: merge G gathered partials [G,4,B,h,w] -> merged [4,B,h,w] = (m, Z, depth = A/Z, eidx = I/Z).
 */
int rmvd_softargmax_combine(const float* partials, int G, int B, int h, int w, float* merged, void* stream);
/*
 * Step 3: this rank's share of the 4-bin confidence around the global index (mvsnet.py:143-158):
 *   conf_part [B,h,w] = sum over d in {idx-1..idx+2} intersect [d_begin,d_end) of exp(x_d - m)/Z, idx = trunc(eidx).
 * The ranks' conf_part are then summed (all-reduce).
 */
int rmvd_softargmax_confidence(const float* cost_slab, const float* merged, int B, int Dl, int h, int w,
                               int D, int d_begin, float* conf_part, void* stream);

/*
 * rmvd_mvs_regress_fwd with D sharded over the num_peers GPUs of one NVLink / NVSwitch box, as ONE kernel per rank (SURVEY.md 8e,
 * partition 1b; replaces rmvd_softargmax_partial + all-gather + rmvd_softargmax_combine + rmvd_softargmax_confidence +
 * all-reduce).  Rank g holds planes [lo_g, lo_g + Dl_g) of the regularised cost (balanced contiguous partition: the first
 * D % num_peers ranks hold one plane more) in a peer-mapped buffer cost_peers[g] laid out [B, Dl_max, h*w], Dl_max =
 * ceil(D / num_peers).  The kernel computes the (m, Z, A, I) partials of its slab (mvsnet.py:140-141 as an online softmax),
 * stores them into slot `rank` of every rank's parts_peers[g] ([num_peers, 4, B*h*w] floats), meets the other ranks on
 * flag_peers[g] (uint32 [2 * num_peers], zero before the first call; `epoch` increases by one per call), merges, and evaluates the
 * 4-bin confidence (mvsnet.py:143-160) reading the four cost values from the ranks that own them.  A second handshake ends the
 * kernel: when it has completed, no rank reads this rank's cost slab or partial buffer any more.  Every rank obtains the full
 * depth / confidence / depth_index / expected_index (each optional, [B,h,w]).  depths: [B, D] on every rank.  local_counters:
 * device uint32 [2], zero.  All ranks must issue the call.
 */
int rmvd_softargmax_slab_fused(const float* const* cost_peers, const float* depths, int B, int D, int h, int w, int num_peers,
                               int rank, float* const* parts_peers, uint32_t* const* flag_peers, uint32_t epoch,
                               uint32_t* local_counters, float* depth, float* confidence, int64_t* depth_index,
                               float* expected_index, void* stream);

/*
 * soft_argmin / entropy of Vis-MVSNet (blocks/utils.py:51-68) along dim 1: volume [B,D,h,w] ->
 * prob [B,D,h,w] (optional), expected index [B,h,w], prob_map [B,h,w] (optional; mass within |d - E| <= window,
 * window < 0 disables), entropy [B,h,w] of the probabilities (optional).
 */
int rmvd_soft_argmin_fwd(const float* volume, int B, int D, int h, int w, float window, float* prob,
                         float* expected_index, float* prob_map, float* entropy, void* stream);

/*
 * Plain reductions over D without a softmax: depth_regression (blocks/utils.py:271-274, `sum_d p[b,d,y,x] * values`) with
 * values per plane [B,D] (mode RMVD_REDUCE_PLANE_VALUES, mvsnet.py:141) or per pixel [B,D,h,w] (RMVD_REDUCE_PIXEL_VALUES,
 * cvp_mvsnet.py:168), and entropy (blocks/utils.py:67-68, `sum_d -v * log(clamp(v, 1e-9, 1))`, RMVD_REDUCE_ENTROPY, values unused).
 * volume [B,D,h,w] -> out [B,h,w].
 */
#define RMVD_REDUCE_PLANE_VALUES 0
#define RMVD_REDUCE_PIXEL_VALUES 1
#define RMVD_REDUCE_ENTROPY 2
int rmvd_volume_reduce_fwd(const float* volume, const float* values, int B, int D, int h, int w, int mode, float* out,
                           void* stream);

/*
 * CVP-MVSNet proj_cost (cvp_mvsnet_components.py:375-457): variance cost volume over PER-PIXEL depth
 * hypotheses, the warped source volumes never written.  feats [B,Vt,C,h,w] (view 0 = reference level features),
 * proj [B,Vt,4,4]: view 0 = inverse(ref_proj), views v >= 1 = src_proj (:398-405), depth_hypos [B,D,h,w],
 * cost [B,C,D,h,w].  Reproduces the reference's in-place quirk (:392): both running sums start from ref^2.
 * rmvd_cvp_projections builds that proj tensor on the device from ref_in [B,3,3], src_in [B,nsrc,3,3], ref_ex [B,4,4],
 * src_ex [B,nsrc,4,4] (:398-405; the reference inverts ref_proj with torch.inverse once per source view).
 */
int rmvd_cvp_projections(const float* ref_in, const float* src_in, const float* ref_ex, const float* src_ex, int B, int nsrc,
                         float* proj, void* stream);
int rmvd_cvp_proj_cost_fwd(const float* feats, const float* proj, const float* depth_hypos, int B, int Vt, int C, int h,
                           int w, int D, float* cost, void* stream);

/* ---------------------------------------------------------------------------------------------------
 * Vis-MVSNet family
 * ------------------------------------------------------------------------------------------------- */

/*
 * build_cost_volume + groupwise_correlation (vis_mvsnet_singlestage.py:86-122, :230-242; blocks/utils.py:71-89, :155-189)
 * fused: cost[n,g,d,y,x] = sum over the channels of group g of ref[n,c,y,x] * homography_warping(src, H[n,d])[n,c,y,x];
 * the warped volume is never written.  H holds one 3x3 homography (9 floats, row-major) per (n, d) or per (n, d, y, x):
 * element (n,d,y,x) starts at H + n*h_stride[0] + d*h_stride[1] + y*h_stride[2] + x*h_stride[3] (0 for broadcast dims),
 * i.e. the output of get_homographies (blocks/utils.py:96-150) as it lies.  ref, src [N,C,h,w]; cost [N,groups,D,h,w].
 */
int rmvd_homography_groupcorr_fwd(const float* ref, const float* src, const float* H, const int64_t* h_stride_host, int N,
                                  int C, int h, int w, int D, int groups, float* cost, void* stream);
/*
 * The same cost volume straight from the cameras: get_homographies (blocks/utils.py:96-150) folded into the kernel, no
 * [n,d,h,w,3,3] tensor.  left_cam, right_cam [N,2,4,4] ([:,0] extrinsics [R|t], [:,1] intrinsics); depth_start and
 * depth_interval are [N,1,1,1] or [N,1,h,w]: element (n,y,x) at n*stride[0] + y*stride[1] + x*stride[2] (0 = broadcast);
 * plane k is depth_start + k*depth_interval, or the inverse-depth schedule of :117-123 when inv != 0.
 * workspace: N*18 floats of device scratch (per-sample A and B with H = A - B / (depth + 1e-9)).
 */
int rmvd_vis_cost_volume_fwd(const float* ref, const float* src, const float* left_cam, const float* right_cam,
                             const float* depth_start, const int64_t* start_stride_host, const float* depth_interval,
                             const int64_t* interval_stride_host, int N, int C, int h, int w, int D, int inv, int groups,
                             float* cost, void* workspace, size_t workspace_bytes, void* stream);
/* homography_warping alone (blocks/utils.py:179-189): warped [N,C,D,h,w] (D = 1 for a single homography per sample). */
int rmvd_homography_warp_fwd(const float* src, const float* H, const int64_t* h_stride_host, int N, int C, int h, int w,
                             int D, float* warped, void* stream);

/*
 * Learned view fusion, elementwise part (rmvd/models/blocks/learned_fusion.py:33-47): replaces
 * softmax(stack(view_weights), 0) + 1e-9, view_weight * mask, the two stacks / products / sums over the views, the division
 * and the fused mask.  logits[v]: [N,1,h,w] (output of corr_to_view_weight, :28), corrs[v] / masks[v]: [N,S,h,w] (the
 * outputs of rmvd_psweep_corr_fwd, separate or views of one allocation); host arrays of V <= 64 device pointers.
 * weights (optional): [V,N,1,h,w] = self.view_weights; fused_corr, fused_mask: [N,S,h,w].  V == 1 is the caller's
 * pass-through (:50-52), not a kernel call.
 */
int rmvd_view_fusion_fwd(const float* const* logits, const float* const* corrs, const float* const* masks, int V, int N,
                         int S, int h, int w, float* weights, float* fused_corr, float* fused_mask, void* stream);

/*
 * Test hook: the sampling chain of the tcgen05 kernel shares one refined reciprocal between its divisions
 * (csrc/common.cuh: rcp_refined / div_shared, the fast path of div.rn.f32 itself).  q_shared[i] = that sequence,
 * q_rn[i] = __fdiv_rn(a[i], b[i]); the two must agree bit for bit for 2^-60 <= |b| <= 2^60, |a| <= 2^60.
 */
int rmvd_debug_div_check(const float* a, const float* b, int n, float* q_shared, float* q_rn, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* RMVD_B200_H_ */
