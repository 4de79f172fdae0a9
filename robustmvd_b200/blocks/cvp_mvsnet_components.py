"""Drop-in for the cost-volume functions of ``rmvd/models/blocks/cvp_mvsnet_components.py`` on top of the sm_100a C ABI
(SURVEY.md row 8f-3): ``homo_warping`` (:192-245) and ``proj_cost`` (:375-457).  Same signatures and return values; the
projection set-up (two 3x4 products and one 4x4 inverse per view) and the per-voxel work both run in the CUDA library.  Forward only, no fallback."""
import torch

from .. import _lib


def _f32c(t):
    return t.detach().to(torch.float32).contiguous()


def _projections(ref_in, src_in, ref_ex, src_ex):
    """:204-211 / :398-405 on the device (``rmvd_cvp_projections``): ``src_in / src_ex [B,nsrc,...]`` ->
    ``proj [B,1+nsrc,4,4]`` with view 0 = ``inverse(ref_proj)`` and view v = ``src_proj`` of source v-1."""
    ref_in, src_in, ref_ex, src_ex = map(_f32c, (ref_in, src_in, ref_ex, src_ex))
    B, nsrc = src_in.shape[0], src_in.shape[1]
    proj = torch.empty((B, 1 + nsrc, 4, 4), dtype=torch.float32, device=ref_in.device)
    st = _lib.load().rmvd_cvp_projections(_lib.ptr(ref_in), _lib.ptr(src_in), _lib.ptr(ref_ex), _lib.ptr(src_ex), B, nsrc,
                                          _lib.ptr(proj), _lib.stream_ptr())
    _lib.check(st, "rmvd_cvp_projections")
    return proj


def homo_warping(src_feature, ref_in, src_in, ref_ex, src_ex, depth_hypos):
    """:192-245: ``src_feature [B,C,H,W]``, intrinsics ``[B,3,3]``, extrinsics ``[B,4,4]``, ``depth_hypos [B,D]`` ->
    warped ``[B,C,D,H,W]`` (``rmvd_homo_warp_fwd``: the chain is MVSNet's ``homo_warp`` up to the order of one product)."""
    proj = _projections(ref_in, src_in[:, None], ref_ex, src_ex[:, None])
    src_proj, ref_proj_inv = proj[:, 1].contiguous(), proj[:, 0].contiguous()
    src_feature, depth_hypos = _f32c(src_feature), _f32c(depth_hypos)
    B, C, H, W = src_feature.shape
    D = depth_hypos.shape[1]
    out = torch.empty((B, C, D, H, W), dtype=torch.float32, device=src_feature.device)
    st = _lib.load().rmvd_homo_warp_fwd(_lib.ptr(src_feature), _lib.ptr(_f32c(src_proj)), _lib.ptr(_f32c(ref_proj_inv)),
                                        _lib.ptr(depth_hypos), B, C, H, W, D, _lib.ptr(out), None, None, _lib.stream_ptr())
    _lib.check(st, "rmvd_homo_warp_fwd")
    return out


def proj_cost(settings, ref_feature, src_feature, level, ref_in, src_in, ref_ex, src_ex, depth_hypos):
    """:375-457: variance cost volume ``[B,C,D,H,W]`` for per-pixel hypotheses ``depth_hypos [B,D,H,W]``;
    ``src_feature[v][level]`` are the source maps, ``src_in / src_ex [B,nsrc,...]``.  One fused launch
    (``rmvd_cvp_proj_cost_fwd``) instead of nsrc warped volumes and ~10 full-volume passes."""
    B, C, H, W = ref_feature.shape
    D = depth_hypos.shape[1]
    nsrc = settings.nsrc
    proj = _projections(ref_in, src_in[:, :nsrc], ref_ex, src_ex[:, :nsrc])                 # [B,1+nsrc,4,4]
    feats = _f32c(torch.stack([ref_feature] + [src_feature[v][level] for v in range(nsrc)], 1))
    cost = torch.empty((B, C, D, H, W), dtype=torch.float32, device=feats.device)
    st = _lib.load().rmvd_cvp_proj_cost_fwd(_lib.ptr(feats), _lib.ptr(proj), _lib.ptr(_f32c(depth_hypos)), B, 1 + nsrc, C, H,
                                            W, D, _lib.ptr(cost), _lib.stream_ptr())
    _lib.check(st, "rmvd_cvp_proj_cost_fwd")
    return cost
