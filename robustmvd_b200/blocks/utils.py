"""Drop-in for the hot-path functions of ``rmvd/models/blocks/utils.py`` and the in-model cost-volume code of
``rmvd/models/mvsnet.py`` on top of the sm_100a C ABI (``include/rmvd_b200.h``).  No fallback; forward only except
``mvs_variance_volume`` / ``mvs_regress`` (the two steps MVSNet trains through).

``homo_warp`` (utils.py:222-268), ``depth_regression`` (:271-274), ``soft_argmin`` (:51-64), ``entropy`` (:67-68)
keep the reference signatures; ``mvs_variance_volume`` and ``mvs_regress`` are the fused forms of
mvsnet.py:123-136 and :140-160 that the MVSNet model calls instead of materialising warped volumes.
"""
import torch

from .. import _lib


def _f32c(t):
    return t.detach().to(torch.float32).contiguous()


def _forward_only(name, *tensors):
    """These wrappers have no autograd node (the reference differentiates through torch ops): refuse, loudly, to return a tensor
    without a graph to a caller that needs gradients (the correlation block and LearnedFusion do have differentiable paths)."""
    if torch.is_grad_enabled() and any(isinstance(t, torch.Tensor) and t.requires_grad for t in tensors):
        raise _lib.RmvdError(f"{name}: forward only -- an input requires grad; call it under torch.no_grad() or detach the inputs")


def homo_warp(src_feat, src_proj, ref_proj_inv, depth_values, return_indices=False):
    """utils.py:222-268: ``src_feat [B,C,H,W]``, projections ``[B,4,4]``, ``depth_values [B,D]`` -> ``[B,C,D,H,W]``."""
    _forward_only("homo_warp", src_feat, src_proj, ref_proj_inv, depth_values)
    src_feat, src_proj, ref_proj_inv, depth_values = map(_f32c, (src_feat, src_proj, ref_proj_inv, depth_values))
    B, C, H, W = src_feat.shape
    D = depth_values.shape[1]
    dev = src_feat.device
    out = torch.empty((B, C, D, H, W), dtype=torch.float32, device=dev)
    ix0 = iy0 = None
    if return_indices:
        ix0 = torch.empty((B, D, H, W), dtype=torch.int32, device=dev)
        iy0 = torch.empty_like(ix0)
    st = _lib.load().rmvd_homo_warp_fwd(_lib.ptr(src_feat), _lib.ptr(src_proj), _lib.ptr(ref_proj_inv),
                                        _lib.ptr(depth_values), B, C, H, W, D, _lib.ptr(out), _lib.ptr(ix0),
                                        _lib.ptr(iy0), _lib.stream_ptr())
    _lib.check(st, "rmvd_homo_warp_fwd")
    return (out, ix0, iy0) if return_indices else out


class _VarianceFunction(torch.autograd.Function):
    """Differentiable ``mvs_variance_volume`` with respect to the feature maps (rmvd_mvs_variance_bwd): what autograd does in the
    reference through homo_warp's grid_sample and the variance expression (mvsnet.py:123-136)."""

    @staticmethod
    def forward(ctx, feats, proj_mats, depth_values):
        with torch.no_grad():
            out = mvs_variance_volume(feats, proj_mats, depth_values)
        ctx.save_for_backward(feats, proj_mats, depth_values)
        return out

    @staticmethod
    def backward(ctx, grad_out):
        feats, proj_mats, depth_values = ctx.saved_tensors
        f, pm, dv, g = map(_f32c, (feats, proj_mats, depth_values, grad_out))
        B, V, C, h, w = f.shape
        grad = torch.empty_like(f)
        st = _lib.load().rmvd_mvs_variance_bwd(_lib.ptr(f), _lib.ptr(pm), _lib.ptr(dv), _lib.ptr(g), B, V, C, h, w, dv.shape[1],
                                               _lib.ptr(grad), _lib.stream_ptr())
        _lib.check(st, "rmvd_mvs_variance_bwd")
        return grad, None, None


class _RegressFunction(torch.autograd.Function):
    """Differentiable expected depth of ``mvs_regress`` with respect to the regularised cost (rmvd_mvs_regress_bwd); confidence and
    indices carry no gradient (the reference computes them under no_grad, mvsnet.py:143-158)."""

    @staticmethod
    def forward(ctx, cost_reg, depth_values):
        with torch.no_grad():
            out = mvs_regress(cost_reg, depth_values)
        ctx.save_for_backward(cost_reg, depth_values)
        ctx.mark_non_differentiable(out["confidence"], out["depth_index"], out["expected_index"])
        return out["depth"], out["confidence"], out["depth_index"], out["expected_index"]

    @staticmethod
    def backward(ctx, grad_depth, *_):
        cost_reg, depth_values = ctx.saved_tensors
        c, dv, g = map(_f32c, (cost_reg, depth_values, grad_depth))
        B, D, h, w = c.shape
        grad = torch.empty_like(c)
        st = _lib.load().rmvd_mvs_regress_bwd(_lib.ptr(c), _lib.ptr(dv), _lib.ptr(g), B, D, h, w, _lib.ptr(grad), _lib.stream_ptr())
        _lib.check(st, "rmvd_mvs_regress_bwd")
        return grad, None


def _needs_grad(*tensors):
    return torch.is_grad_enabled() and any(isinstance(t, torch.Tensor) and t.requires_grad for t in tensors)


def mvs_variance_volume(feats, proj_mats, depth_values, plane_range=None):
    """mvsnet.py:123-136 fused with homo_warp: ``feats [B,V,C,h,w]`` (view 0 = key), ``proj_mats [B,V,4,4]`` (view 0
    already inverted), ``depth_values [B,D]`` -> variance ``[B,C,D,h,w]`` (or the ``plane_range`` slab of D).
    Differentiable with respect to ``feats`` (full plane range only)."""
    if _needs_grad(feats) and plane_range is None and not _needs_grad(proj_mats, depth_values):
        return _VarianceFunction.apply(feats, proj_mats, depth_values)
    _forward_only("mvs_variance_volume", feats, proj_mats, depth_values)
    feats, proj_mats, depth_values = map(_f32c, (feats, proj_mats, depth_values))
    B, V, C, h, w = feats.shape
    D = depth_values.shape[1]
    d0, d1 = (0, D) if plane_range is None else plane_range
    out = torch.empty((B, C, d1 - d0, h, w), dtype=torch.float32, device=feats.device)
    lib = _lib.load()
    need = lib.rmvd_mvs_variance_workspace_bytes(B, V, C, h, w)     # channels-last copy of the source views (0: shape takes the NCHW kernel)
    ws = torch.empty(max(need, 16), dtype=torch.uint8, device=feats.device)
    st = lib.rmvd_mvs_variance_fwd_ws(_lib.ptr(feats), _lib.ptr(proj_mats), _lib.ptr(depth_values), B, V, C, h, w, D, d0, d1,
                                      _lib.ptr(out), d1 - d0, 0, _lib.ptr(ws), need, _lib.stream_ptr())
    _lib.check(st, "rmvd_mvs_variance_fwd_ws")
    return out


def mvs_regress(cost_reg, depth_values, return_prob=False):
    """mvsnet.py:140-160 in one pass: ``cost_reg [B,D,h,w]`` (pre-softmax), ``depth_values [B,D]`` ->
    dict(depth, confidence, depth_index (int64), expected_index[, prob]).  ``depth`` is differentiable with respect to ``cost_reg``."""
    if _needs_grad(cost_reg) and not return_prob and not _needs_grad(depth_values):
        depth, conf, didx, eidx = _RegressFunction.apply(cost_reg, depth_values)
        return dict(depth=depth, confidence=conf, depth_index=didx, expected_index=eidx)
    _forward_only("mvs_regress", cost_reg, depth_values)
    cost_reg, depth_values = _f32c(cost_reg), _f32c(depth_values)
    B, D, h, w = cost_reg.shape
    dev = cost_reg.device
    depth = torch.empty((B, h, w), dtype=torch.float32, device=dev)
    conf = torch.empty_like(depth)
    eidx = torch.empty_like(depth)
    didx = torch.empty((B, h, w), dtype=torch.int64, device=dev)
    prob = torch.empty_like(cost_reg) if return_prob else None
    st = _lib.load().rmvd_mvs_regress_fwd(_lib.ptr(cost_reg), _lib.ptr(depth_values), B, D, h, w, _lib.ptr(depth),
                                          _lib.ptr(conf), _lib.ptr(didx), _lib.ptr(eidx), _lib.ptr(prob),
                                          _lib.stream_ptr())
    _lib.check(st, "rmvd_mvs_regress_fwd")
    out = dict(depth=depth, confidence=conf, depth_index=didx, expected_index=eidx)
    if return_prob:
        out["prob"] = prob
    return out


def _volume_reduce(volume, values, mode):
    volume = _f32c(volume)
    B, D, h, w = volume.shape
    out = torch.empty((B, h, w), dtype=torch.float32, device=volume.device)
    st = _lib.load().rmvd_volume_reduce_fwd(_lib.ptr(volume), _lib.ptr(values), B, D, h, w, mode, _lib.ptr(out), _lib.stream_ptr())
    _lib.check(st, "rmvd_volume_reduce_fwd")
    return out


def depth_regression(p, depth_values):
    """utils.py:271-274: ``sum_d p[:,d] * depth_values[:,d]`` for ``p [B,D,h,w]`` (normalised or not: no softmax is applied) and
    ``depth_values [B,D]`` (mvsnet.py:141), ``[D]`` or per-pixel ``[B,D,h,w]`` (cvp_mvsnet.py:168)."""
    _forward_only("depth_regression", p, depth_values)
    if depth_values.ndim == 1:
        depth_values = depth_values[None].expand(p.shape[0], -1)
    if depth_values.ndim == 4:
        assert tuple(depth_values.shape) == tuple(p.shape)
        return _volume_reduce(p, _f32c(depth_values), 1)
    assert depth_values.ndim == 2 and depth_values.shape[1] == p.shape[1]
    return _volume_reduce(p, _f32c(depth_values.expand(p.shape[0], -1)), 0)


def soft_argmin(volume, dim, keepdim=False, window=None):
    """utils.py:51-64 along ``dim`` = 1 of ``[B,D,h,w]``."""
    _forward_only("soft_argmin", volume)
    assert dim == 1 and volume.ndim == 4, "the kernel reduces dim 1 of a [B,D,h,w] volume"
    volume = _f32c(volume)
    B, D, h, w = volume.shape
    prob = torch.empty_like(volume)
    eidx = torch.empty((B, h, w), dtype=torch.float32, device=volume.device)
    pmap = torch.empty_like(eidx) if window is not None else None
    st = _lib.load().rmvd_soft_argmin_fwd(_lib.ptr(volume), B, D, h, w, -1.0 if window is None else float(window),
                                          _lib.ptr(prob), _lib.ptr(eidx), _lib.ptr(pmap), None, _lib.stream_ptr())
    _lib.check(st, "rmvd_soft_argmin_fwd")
    out = eidx.unsqueeze(1) if keepdim else eidx
    if window is None:
        return prob, out
    return prob, out, (pmap.unsqueeze(1) if keepdim else pmap)


def entropy(volume, dim, keepdim=False):
    """utils.py:67-68: ``sum_d -v * log(clamp(v, 1e-9, 1))`` along dim 1 of ``[B,D,h,w]`` (no normalisation, as in the reference)."""
    assert dim == 1 and volume.ndim == 4
    _forward_only("entropy", volume)
    ent = _volume_reduce(volume, None, 2)
    return ent.unsqueeze(1) if keepdim else ent


# ---------------------------------------------------------------------------------------------------------
# depth-slab softargmax: the three local steps of the multi-GPU regression
# ---------------------------------------------------------------------------------------------------------
def softargmax_partial(cost_slab, depth_slab, d_begin):
    """Online-softmax partials of a slab ``cost_slab [B,Dl,h,w]`` holding planes ``[d_begin, d_begin + Dl)``:
    ``[4,B,h,w] = (m, Z = sum exp(x-m), A = sum exp(x-m)*depth, I = sum exp(x-m)*index)``."""
    cost_slab, depth_slab = _f32c(cost_slab), _f32c(depth_slab)
    B, Dl, h, w = cost_slab.shape
    out = torch.empty((4, B, h, w), dtype=torch.float32, device=cost_slab.device)
    st = _lib.load().rmvd_softargmax_partial(_lib.ptr(cost_slab), _lib.ptr(depth_slab), B, Dl, h, w, int(d_begin),
                                             _lib.ptr(out), _lib.stream_ptr())
    _lib.check(st, "rmvd_softargmax_partial")
    return out


def softargmax_combine(partials):
    """Merge gathered partials ``[G,4,B,h,w]`` -> ``[4,B,h,w] = (m, Z, depth, expected index)``."""
    partials = _f32c(partials)
    G, _, B, h, w = partials.shape
    out = torch.empty((4, B, h, w), dtype=torch.float32, device=partials.device)
    st = _lib.load().rmvd_softargmax_combine(_lib.ptr(partials), G, B, h, w, _lib.ptr(out), _lib.stream_ptr())
    _lib.check(st, "rmvd_softargmax_combine")
    return out


def softargmax_confidence(cost_slab, merged, D, d_begin):
    """This slab's share of the 4-bin confidence around the global index (mvsnet.py:143-158): ``[B,h,w]``."""
    cost_slab, merged = _f32c(cost_slab), _f32c(merged)
    B, Dl, h, w = cost_slab.shape
    out = torch.empty((B, h, w), dtype=torch.float32, device=cost_slab.device)
    st = _lib.load().rmvd_softargmax_confidence(_lib.ptr(cost_slab), _lib.ptr(merged), B, Dl, h, w, int(D),
                                                int(d_begin), _lib.ptr(out), _lib.stream_ptr())
    _lib.check(st, "rmvd_softargmax_confidence")
    return out


# ---------------------------------------------------------------------------------------------------------------------
# Vis-MVSNet cost volume: rmvd/models/blocks/utils.py:71-189, vis_mvsnet_singlestage.py:86-122, :230-242
# ---------------------------------------------------------------------------------------------------------------------
class NanError(Exception):
    """utils.py:91-92."""


def get_homographies(left_cam, right_cam, depth_num, depth_start, depth_interval, inv=False):
    """utils.py:96-150: plane-induced homographies ``H = K_r R_r (I - (c_r - c_l) n^T / depth) R_l^T K_l^-1`` for
    ``depth_num`` fronto-parallel planes of the left (reference) camera.  ``left_cam, right_cam [n,2,4,4]`` (``[:,0]``
    extrinsics ``[R|t]``, ``[:,1]`` intrinsics), ``depth_start, depth_interval [n,1,1,1]`` or ``[n,1,h,w]``.  Returns
    ``[n, depth_num, 1|h, 1|w, 3, 3]``.  Host-side set-up like the reference's (a handful of 3x3 products per plane, or per
    pixel in the refinement stages); the per-voxel work is in the CUDA kernels that consume the result."""
    with torch.no_grad():
        n = depth_start.shape[0]
        dt, dev = left_cam.dtype, left_cam.device
        steps = torch.arange(0, depth_num, dtype=dt, device=dev).view(1, depth_num, 1, 1)
        if not inv:
            depth = depth_start + depth_interval * steps
        else:
            depth_end = depth_start + (depth_num - 1) * depth_interval
            inv_interv = (1 / (depth_start + 1e-9) - 1 / (depth_end + 1e-9)) / (depth_num - 1 + 1e-9)
            depth = 1 / (1 / (depth_end + 1e-9) + inv_interv * steps)
        depth = depth.unsqueeze(-1).unsqueeze(-1)                                    # n d h w 1 1
        R_l, R_r = left_cam[:, 0, :3, :3], right_cam[:, 0, :3, :3]
        t_l, t_r = left_cam[:, 0, :3, 3:4], right_cam[:, 0, :3, 3:4]
        K_l, K_r = left_cam[:, 1, :3, :3], right_cam[:, 1, :3, :3]
        K_l_inv = K_l.float().inverse().to(dt)
        R_l_t, R_r_t = R_l.transpose(-2, -1), R_r.transpose(-2, -1)
        fronto = R_l[:, 2:3, :3]                                                     # n 1 3
        c_rel = (-R_r_t @ t_r) - (-R_l_t @ t_l)                                      # n 3 1
        outer = (c_rel @ fronto).view(n, 1, 1, 1, 3, 3)
        middle0 = torch.eye(3, dtype=dt, device=dev).view(1, 1, 1, 1, 3, 3) - outer / (depth + 1e-9)
        middle1 = (R_l_t @ K_l_inv).view(n, 1, 1, 1, 3, 3)
        homographies = K_r.view(n, 1, 1, 1, 3, 3) @ R_r.view(n, 1, 1, 1, 3, 3) @ (middle0 @ middle1)
    if (homographies != homographies).any():
        raise NanError
    return homographies


def _h_strides(Hs):
    """Element strides (n, d, y, x) of contiguous homographies ``[n,d,1|h,1|w,3,3]``, broadcast dims -> 0."""
    st = list(Hs.stride()[:4])
    for k in (2, 3):
        if Hs.shape[k] == 1:
            st[k] = 0
    return st


def homography_warping(input, H):
    """utils.py:179-189: ``input [n,c,h,w]``, ``H [n,3,3]`` or ``[n,h,w,3,3]`` -> warped ``[n,c,h,w]``
    (``rmvd_homography_warp_fwd`` with one plane)."""
    input, H = _f32c(input), _f32c(H)
    n, c, h, w = input.shape
    Hs = H.view(n, 1, 1, 1, 3, 3) if H.ndim == 3 else H.view(n, 1, h, w, 3, 3)
    out = torch.empty((n, c, 1, h, w), dtype=torch.float32, device=input.device)
    st = _lib.load().rmvd_homography_warp_fwd(_lib.ptr(input), _lib.ptr(Hs), _lib.i64_array(_h_strides(Hs)), n, c, h, w, 1,
                                              _lib.ptr(out), _lib.stream_ptr())
    _lib.check(st, "rmvd_homography_warp_fwd")
    return out[:, :, 0]


def homography_warp_volume(src, Hs):
    """``build_cost_volume`` (vis_mvsnet_singlestage.py:86-122) from ready homographies ``[n,d,1|h,1|w,3,3]``:
    warped ``[n,c,d,h,w]`` without the d-fold repeat of the source (:109-113)."""
    src, Hs = _f32c(src), _f32c(Hs)
    n, c, h, w = src.shape
    d = Hs.shape[1]
    out = torch.empty((n, c, d, h, w), dtype=torch.float32, device=src.device)
    st = _lib.load().rmvd_homography_warp_fwd(_lib.ptr(src), _lib.ptr(Hs), _lib.i64_array(_h_strides(Hs)), n, c, h, w, d,
                                              _lib.ptr(out), _lib.stream_ptr())
    _lib.check(st, "rmvd_homography_warp_fwd")
    return out


def groupwise_cost_volume(ref, src, Hs, groups=8):
    """``groupwise_correlation(ref_ncdhw, build_cost_volume(...), groups, 1)`` (vis_mvsnet_singlestage.py:230-242,
    utils.py:71-89) in one pass: ``ref, src [n,c,h,w]``, ``Hs [n,d,1|h,1|w,3,3]`` -> ``[n,groups,d,h,w]``; the warped
    volume is never materialised."""
    ref, src, Hs = _f32c(ref), _f32c(src), _f32c(Hs)
    n, c, h, w = src.shape
    assert ref.shape == src.shape and c % groups == 0
    d = Hs.shape[1]
    out = torch.empty((n, groups, d, h, w), dtype=torch.float32, device=src.device)
    st = _lib.load().rmvd_homography_groupcorr_fwd(_lib.ptr(ref), _lib.ptr(src), _lib.ptr(Hs),
                                                   _lib.i64_array(_h_strides(Hs)), n, c, h, w, d, groups, _lib.ptr(out),
                                                   _lib.stream_ptr())
    _lib.check(st, "rmvd_homography_groupcorr_fwd")
    return out


def vis_cost_volume(ref, src, left_cam, right_cam, depth_num, depth_start, depth_interval, groups=8, inv=False):
    """``get_homographies`` + ``build_cost_volume`` + ``groupwise_correlation`` (utils.py:96-150, vis_mvsnet_singlestage.py:
    86-122, :230-242) in one call: the homography of every (plane, pixel) is formed inside the kernel from two 3x3 matrices per
    sample, so neither ``Hs [n,d,h,w,3,3]`` nor the warped volume exists.  ``left_cam / right_cam [n,2,4,4]`` (already scaled,
    :99-100), ``depth_start / depth_interval [n,1,1,1]`` or ``[n,1,h,w]`` -> ``[n,groups,depth_num,h,w]``."""
    ref, src, left_cam, right_cam = map(_f32c, (ref, src, left_cam, right_cam))
    ds, di = _f32c(depth_start), _f32c(depth_interval)
    n, c, h, w = src.shape

    def strides(t):
        assert t.shape[0] == n and t.shape[1] == 1 and t.shape[2] in (1, h) and t.shape[3] in (1, w)
        st = t.stride()
        return [st[0], 0 if t.shape[2] == 1 else st[2], 0 if t.shape[3] == 1 else st[3]]

    out = torch.empty((n, groups, depth_num, h, w), dtype=torch.float32, device=src.device)
    wsp = torch.empty(n * 18, dtype=torch.float32, device=src.device)
    st = _lib.load().rmvd_vis_cost_volume_fwd(
        _lib.ptr(ref), _lib.ptr(src), _lib.ptr(left_cam), _lib.ptr(right_cam), _lib.ptr(ds), _lib.i64_array(strides(ds)),
        _lib.ptr(di), _lib.i64_array(strides(di)), n, c, h, w, depth_num, 1 if inv else 0, groups, _lib.ptr(out),
        _lib.ptr(wsp), wsp.numel() * 4, _lib.stream_ptr())
    _lib.check(st, "rmvd_vis_cost_volume_fwd")
    return out
