"""Drop-in for the cost-volume functions of ``rmvd/models/blocks/cvp_mvsnet_components.py`` on top of the sm_100a C ABI
(SURVEY.md row 8f-3): ``homo_warping`` (:192-245) and ``proj_cost`` (:375-457).  Same signatures and return values; the
projection set-up (two 3x4 products and one 4x4 inverse per view) stays a host-side torch expression like in the reference,
the per-voxel work runs in the CUDA library.  Forward only, no fallback."""
import torch

from .. import _lib


def _f32c(t):
    return t.detach().to(torch.float32).contiguous()


def _proj4(K, E):
    """:204-209 / :398-403: ``K @ E[:3]`` with the row (0, 0, 0, 1) appended (built on the device, no host copy)."""
    out = torch.zeros((K.shape[0], 4, 4), device=K.device, dtype=K.dtype)
    out[:, :3] = torch.matmul(K, E[:, 0:3, :])
    out[:, 3, 3] = 1.0
    return out


def _projections(ref_in, src_in, ref_ex, src_ex, ref_proj_inv=None):
    """:204-211 / :398-405: ``src_proj`` and ``inverse(ref_proj)`` as 4x4 matrices (``inv_ex`` without the error check: no
    device-to-host synchronisation; the reference inverts the same reference projection once per source view)."""
    with torch.no_grad():
        if ref_proj_inv is None:
            ref_proj_inv = torch.linalg.inv_ex(_proj4(ref_in, ref_ex), check_errors=False)[0]
        return _proj4(src_in, src_ex), ref_proj_inv


def homo_warping(src_feature, ref_in, src_in, ref_ex, src_ex, depth_hypos):
    """:192-245: ``src_feature [B,C,H,W]``, intrinsics ``[B,3,3]``, extrinsics ``[B,4,4]``, ``depth_hypos [B,D]`` ->
    warped ``[B,C,D,H,W]`` (``rmvd_homo_warp_fwd``: the chain is MVSNet's ``homo_warp`` up to the order of one product)."""
    src_proj, ref_proj_inv = _projections(ref_in, src_in, ref_ex, src_ex)
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
    projs, ref_proj_inv = [], None
    for v in range(nsrc):
        src_proj, ref_proj_inv = _projections(ref_in, src_in[:, v], ref_ex, src_ex[:, v], ref_proj_inv)
        if v == 0:
            projs.append(ref_proj_inv)
        projs.append(src_proj)
    proj = _f32c(torch.stack(projs, 1))                                                    # [B,1+nsrc,4,4]
    feats = _f32c(torch.stack([ref_feature] + [src_feature[v][level] for v in range(nsrc)], 1))
    cost = torch.empty((B, C, D, H, W), dtype=torch.float32, device=feats.device)
    st = _lib.load().rmvd_cvp_proj_cost_fwd(_lib.ptr(feats), _lib.ptr(proj), _lib.ptr(_f32c(depth_hypos)), B, 1 + nsrc, C, H,
                                            W, D, _lib.ptr(cost), _lib.stream_ptr())
    _lib.check(st, "rmvd_cvp_proj_cost_fwd")
    return cost
