"""Drop-in for the hot-path functions of ``rmvd/models/blocks/utils.py`` and the in-model cost-volume code of
``rmvd/models/mvsnet.py`` on top of the sm_100a C ABI (``include/rmvd_b200.h``).  Forward only, no fallback.

``homo_warp`` (utils.py:222-268), ``depth_regression`` (:271-274), ``soft_argmin`` (:51-64), ``entropy`` (:67-68)
keep the reference signatures; ``mvs_variance_volume`` and ``mvs_regress`` are the fused forms of
mvsnet.py:123-136 and :140-160 that the MVSNet model calls instead of materialising warped volumes.
"""
import torch

from .. import _lib


def _f32c(t):
    return t.detach().to(torch.float32).contiguous()


def homo_warp(src_feat, src_proj, ref_proj_inv, depth_values, return_indices=False):
    """utils.py:222-268: ``src_feat [B,C,H,W]``, projections ``[B,4,4]``, ``depth_values [B,D]`` -> ``[B,C,D,H,W]``."""
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


def mvs_variance_volume(feats, proj_mats, depth_values, plane_range=None):
    """mvsnet.py:123-136 fused with homo_warp: ``feats [B,V,C,h,w]`` (view 0 = key), ``proj_mats [B,V,4,4]`` (view 0
    already inverted), ``depth_values [B,D]`` -> variance ``[B,C,D,h,w]`` (or the ``plane_range`` slab of D)."""
    feats, proj_mats, depth_values = map(_f32c, (feats, proj_mats, depth_values))
    B, V, C, h, w = feats.shape
    D = depth_values.shape[1]
    d0, d1 = (0, D) if plane_range is None else plane_range
    out = torch.empty((B, C, d1 - d0, h, w), dtype=torch.float32, device=feats.device)
    st = _lib.load().rmvd_mvs_variance_fwd(_lib.ptr(feats), _lib.ptr(proj_mats), _lib.ptr(depth_values), B, V, C, h, w,
                                           D, d0, d1, _lib.ptr(out), d1 - d0, 0, _lib.stream_ptr())
    _lib.check(st, "rmvd_mvs_variance_fwd")
    return out


def mvs_regress(cost_reg, depth_values, return_prob=False):
    """mvsnet.py:140-160 in one pass: ``cost_reg [B,D,h,w]`` (pre-softmax), ``depth_values [B,D]`` ->
    dict(depth, confidence, depth_index (int64), expected_index[, prob])."""
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


def depth_regression(p, depth_values):
    """utils.py:271-274: ``sum_d p[:,d] * depth_values[:,d]`` for an already normalised ``p [B,D,h,w]``.
    Runs the regression kernel on ``log p`` (softmax(log p) = p for a normalised p)."""
    p = _f32c(p)
    if depth_values.ndim == 1:
        depth_values = depth_values[None].expand(p.shape[0], -1)
    return mvs_regress(torch.log(p.clamp_min(1e-38)), depth_values)["depth"]


def soft_argmin(volume, dim, keepdim=False, window=None):
    """utils.py:51-64 along ``dim`` = 1 of ``[B,D,h,w]``."""
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
    """utils.py:67-68 for a probability volume ``[B,D,h,w]`` along dim 1 (softmax(log p) = p)."""
    assert dim == 1 and volume.ndim == 4
    volume = _f32c(volume)
    B, D, h, w = volume.shape
    eidx = torch.empty((B, h, w), dtype=torch.float32, device=volume.device)
    ent = torch.empty_like(eidx)
    st = _lib.load().rmvd_soft_argmin_fwd(_lib.ptr(torch.log(volume.clamp_min(1e-38))), B, D, h, w, -1.0, None,
                                          _lib.ptr(eidx), None, _lib.ptr(ent), _lib.stream_ptr())
    _lib.check(st, "rmvd_soft_argmin_fwd")
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
