"""Oracle (test infrastructure): numpy restatement of the MVSNet cost-volume path.

Follows ``/root/reference/rmvd/models/blocks/utils.py:222-274`` (``homo_warp``, ``depth_regression``),
``:51-68`` (``soft_argmin``, ``entropy``) and ``/root/reference/rmvd/models/mvsnet.py:123-160``.
Not imported by the product package.
"""
import numpy as np

from .planesweep import F32, _f32, _fma32, grid_unnormalize

_ONE = F32(1.0)


def _matmul4(a, b):
    """Batched 4x4 product in fp32 (utils.py:234): k-ascending, every multiply and add rounded separately
    (what torch 2.11 CPU's small batched matmul does: 100 % bit agreement in tools/validate_oracle.py)."""
    a = _f32(a)
    b = _f32(b)
    out = np.zeros(a.shape[:-2] + (4, 4), np.float32)
    for i in range(4):
        for j in range(4):
            acc = a[..., i, 0] * b[..., 0, j]
            for k in range(1, 4):
                acc = acc + a[..., i, k] * b[..., k, j]
            out[..., i, j] = acc
    return out


def homo_warp_coords(src_proj, ref_proj_inv, depth_values, H, W):
    """utils.py:234-259: source-view pixel coordinates and the ATen-unnormalised sampling coordinates.

    Returns ``ix, iy`` (``[B,D,H,W]`` fp32, already unnormalised: the value whose floor is the corner index).
    The 3x3 product ``R @ (x*d, y*d, d) + T`` (:251) runs through BLAS sgemm in the reference; the
    restatement accumulates k-ascending with fused multiply-adds, which is what the probed torch CPU build
    does (tools/validate_oracle.py reports the index agreement).
    """
    P = _matmul4(src_proj, ref_proj_inv)  # :234
    R = P[:, :3, :3]
    T = P[:, :3, 3]
    depth = _f32(depth_values)
    B, D = depth.shape
    xs = np.arange(W, dtype=np.float32)[None, None, None, :]  # create_meshgrid(normalized=False): 0..W-1
    ys = np.arange(H, dtype=np.float32)[None, None, :, None]
    d = depth[:, :, None, None]
    gx = xs * d  # :248
    gy = ys * d
    gz = _ONE * d * np.ones((1, 1, H, 1), np.float32)
    gz = np.broadcast_to(gz, (B, D, H, 1))

    def row(i):
        r0 = R[:, i, 0][:, None, None, None]
        r1 = R[:, i, 1][:, None, None, None]
        r2 = R[:, i, 2][:, None, None, None]
        t = T[:, i][:, None, None, None]
        acc = r0 * gx
        acc = _fma32(r1, gy, acc)
        acc = _fma32(r2, gz, acc)
        return (acc + t).astype(np.float32)  # :251

    with np.errstate(divide="ignore", invalid="ignore", over="ignore"):
        px, py, pz = row(0), row(1), row(2)
        sx = px / pz  # :253
        sy = py / pz
        xn = sx / F32((W - 1) / 2) - _ONE  # :256
        yn = sy / F32((H - 1) / 2) - _ONE  # :257
        ix = grid_unnormalize(xn, W)  # F.grid_sample align_corners=False (:262-264)
        iy = grid_unnormalize(yn, H)
    return ix, iy


def _bilinear_from_unnormalized(ix, iy, H, W):
    with np.errstate(invalid="ignore", over="ignore"):
        x0 = np.floor(ix)
        y0 = np.floor(iy)
        wx1 = ix - x0
        wx0 = _ONE - wx1
        wy1 = iy - y0
        wy0 = _ONE - wy1
    finite = np.isfinite(ix) & np.isfinite(iy)
    x0_in = (x0 >= 0) & (x0 <= W - 1) & finite
    x1_in = (x0 + 1 >= 0) & (x0 + 1 <= W - 1) & finite
    y0_in = (y0 >= 0) & (y0 <= H - 1) & finite
    y1_in = (y0 + 1 >= 0) & (y0 + 1 <= H - 1) & finite
    with np.errstate(invalid="ignore"):
        ix0 = np.clip(np.nan_to_num(x0, nan=-2.0), -2, W).astype(np.int32)
        iy0 = np.clip(np.nan_to_num(y0, nan=-2.0), -2, H).astype(np.int32)
    return dict(ix0=ix0, iy0=iy0, nw=wy0 * wx0, ne=wy0 * wx1, sw=wy1 * wx0, se=wy1 * wx1,
                in_nw=y0_in & x0_in, in_ne=y0_in & x1_in, in_sw=y1_in & x0_in, in_se=y1_in & x1_in)


def _gather_bilinear(img, b):
    """img ``[C,H,W]``; b fields ``[D,H,W]`` -> ``[C,D,H,W]``; out-of-bounds corners contribute zero."""
    C, H, W = img.shape
    flat = img.reshape(C, H * W)
    x0 = np.clip(b["ix0"], 0, W - 1).astype(np.int64)
    x1 = np.clip(b["ix0"] + 1, 0, W - 1).astype(np.int64)
    y0 = np.clip(b["iy0"], 0, H - 1).astype(np.int64)
    y1 = np.clip(b["iy0"] + 1, 0, H - 1).astype(np.int64)
    z = F32(0)
    nz = lambda wgt: np.nan_to_num(wgt, nan=0.0, posinf=0.0, neginf=0.0)  # noqa: E731
    acc = np.where(b["in_nw"], flat[:, y0 * W + x0], z) * np.where(b["in_nw"], nz(b["nw"]), z)
    acc = acc + np.where(b["in_ne"], flat[:, y0 * W + x1], z) * np.where(b["in_ne"], nz(b["ne"]), z)
    acc = acc + np.where(b["in_sw"], flat[:, y1 * W + x0], z) * np.where(b["in_sw"], nz(b["sw"]), z)
    acc = acc + np.where(b["in_se"], flat[:, y1 * W + x1], z) * np.where(b["in_se"], nz(b["se"]), z)
    return acc.astype(np.float32)


def homo_warp(src_feat, src_proj, ref_proj_inv, depth_values, return_debug=False):
    """utils.py:222-268.  ``src_feat [B,C,H,W]`` -> warped ``[B,C,D,H,W]`` (bilinear, zeros padding, no mask)."""
    src_feat = _f32(src_feat)
    B, C, H, W = src_feat.shape
    ix, iy = homo_warp_coords(src_proj, ref_proj_inv, depth_values, H, W)
    D = ix.shape[1]
    out = np.empty((B, C, D, H, W), np.float32)
    dbg = []
    for bi in range(B):
        b = _bilinear_from_unnormalized(ix[bi], iy[bi], H, W)
        out[bi] = _gather_bilinear(src_feat[bi], b)
        dbg.append(b)
    if return_debug:
        return out, dict(ix0=np.stack([d["ix0"] for d in dbg]), iy0=np.stack([d["iy0"] for d in dbg]))
    return out


def mvs_variance(feats, proj_mats, depth_values):
    """mvsnet.py:123-136.  ``feats [B,V,C,h,w]`` (view 0 = key), ``proj_mats [B,V,4,4]`` (view 0 already
    inverted, :85-86), ``depth_values [B,D]`` -> variance volume ``[B,C,D,h,w]``."""
    feats = _f32(feats)
    proj_mats = _f32(proj_mats)
    B, V, C, h, w = feats.shape
    D = depth_values.shape[1]
    ref = feats[:, 0][:, :, None]  # :124
    vsum = np.broadcast_to(ref, (B, C, D, h, w)).astype(np.float32)  # :125
    vsq = vsum ** 2  # :126
    for v in range(1, V):  # :129-133
        warped = homo_warp(feats[:, v], proj_mats[:, v], proj_mats[:, 0], depth_values)
        vsum = vsum + warped
        vsq = vsq + warped ** 2
    Vf = F32(V)
    mean = vsum / Vf
    return (vsq / Vf - mean * mean).astype(np.float32)  # :135


def mvs_variance_backward(feats, proj_mats, depth_values, grad_variance):
    """Gradient of ``sum <grad_variance, mvs_variance(feats, ...)>`` with respect to ``feats`` -- what autograd returns in the
    reference through homo_warp's grid_sample (utils.py:262) and mvsnet.py:123-136: ``d var / d x_v = (2 / V) (x_v - mean)``;
    a source view receives it through the transpose of its bilinear gather (zeros padding: no gradient from corners outside)."""
    feats = _f32(feats)
    proj_mats = _f32(proj_mats)
    g = _f32(grad_variance)
    B, V, C, h, w = feats.shape
    D = depth_values.shape[1]
    warped = [None] + [homo_warp(feats[:, v], proj_mats[:, v], proj_mats[:, 0], depth_values) for v in range(1, V)]
    ref = np.broadcast_to(feats[:, 0][:, :, None], (B, C, D, h, w))
    mean = (ref + sum(warped[1:])) / F32(V)
    out = np.zeros_like(feats, dtype=np.float64)
    out[:, 0] = (2.0 / V * g * (ref - mean)).sum(axis=2)
    for v in range(1, V):
        gx = (2.0 / V) * g.astype(np.float64) * (warped[v] - mean)      # [B,C,D,h,w]
        ix, iy = homo_warp_coords(proj_mats[:, v], proj_mats[:, 0], depth_values, h, w)
        for bi in range(B):
            b = _bilinear_from_unnormalized(ix[bi], iy[bi], h, w)
            x0 = np.clip(b["ix0"], 0, w - 1).astype(np.int64)
            x1 = np.clip(b["ix0"] + 1, 0, w - 1).astype(np.int64)
            y0 = np.clip(b["iy0"], 0, h - 1).astype(np.int64)
            y1 = np.clip(b["iy0"] + 1, 0, h - 1).astype(np.int64)
            flat = out[bi, v].reshape(C, h * w)
            for key, yy, xx in (("nw", y0, x0), ("ne", y0, x1), ("sw", y1, x0), ("se", y1, x1)):
                wgt = np.where(b["in_" + key], np.nan_to_num(b[key], nan=0.0, posinf=0.0, neginf=0.0), 0.0).astype(np.float64)
                idx = (yy * w + xx).reshape(-1)
                contrib = (gx[bi] * wgt[None]).reshape(C, -1)
                for c in range(C):
                    np.add.at(flat[c], idx, contrib[c])
    return out.astype(np.float32)


def mvs_regress_backward(cost_reg, depth_values, grad_depth):
    """Gradient of ``sum <grad_depth, depth>`` with respect to ``cost_reg`` for ``depth = sum_d softmax(cost)_d * depth_d``
    (mvsnet.py:140-141): ``g * p_d * (depth_d - depth)``."""
    p = _softmax(cost_reg, 1).astype(np.float64)
    B, D, h, w = p.shape
    dv = _f32(depth_values).reshape(B, D, 1, 1).astype(np.float64)
    depth = np.sum(p * dv, axis=1, keepdims=True)
    return (_f32(grad_depth)[:, None].astype(np.float64) * p * (dv - depth)).astype(np.float32)


def _softmax(x, axis):
    x = _f32(x)
    m = x.max(axis=axis, keepdims=True)
    e = np.exp(x - m, dtype=np.float32)
    return (e / e.sum(axis=axis, keepdims=True, dtype=np.float32)).astype(np.float32)


def mvs_regress(cost_reg, depth_values):
    """mvsnet.py:140-160.  ``cost_reg [B,D,h,w]``, ``depth_values [B,D]``.

    Returns ``depth [B,h,w]``, ``confidence [B,h,w]``, ``depth_index [B,h,w] int64`` and the fp32 expected index
    (so tests can exclude pixels whose expected index sits within rounding of an integer).
    """
    p = _softmax(cost_reg, 1)  # :140
    B, D, h, w = p.shape
    dv = _f32(depth_values).reshape(B, D, 1, 1)
    depth = np.sum(p * dv, axis=1, dtype=np.float32)  # :141, utils.py:271-274
    pad = np.zeros((B, D + 3, h, w), np.float32)  # :145-151  pad 1 before, 2 after
    pad[:, 1:D + 1] = p
    p4 = ((pad[:, 0:D] + pad[:, 1:D + 1]) + pad[:, 2:D + 2]) + pad[:, 3:D + 3]
    eidx = np.sum(p * np.arange(D, dtype=np.float32).reshape(1, D, 1, 1), axis=1, dtype=np.float32)  # :153-156
    idx = np.clip(eidx.astype(np.int64), 0, D - 1)
    conf = np.take_along_axis(p4, idx[:, None], axis=1)[:, 0]  # :158
    return depth.astype(np.float32), conf.astype(np.float32), idx, eidx.astype(np.float32)


def softargmax_partials(cost_slab, depth_slab, d_begin):
    """Per-slab online-softmax partials (partition 1b): for planes ``[d_begin, d_begin+Dl)``
    returns ``m = max``, ``Z = sum exp(x-m)``, ``A = sum exp(x-m) * depth``, ``I = sum exp(x-m) * index``,
    each ``[B,h,w]``."""
    x = _f32(cost_slab)
    B, Dl, h, w = x.shape
    m = x.max(axis=1)
    e = np.exp(x - m[:, None], dtype=np.float32)
    Z = e.sum(axis=1, dtype=np.float32)
    A = np.sum(e * _f32(depth_slab).reshape(B, Dl, 1, 1), axis=1, dtype=np.float32)
    idx = (np.arange(Dl, dtype=np.float32) + F32(d_begin)).reshape(1, Dl, 1, 1)
    I = np.sum(e * idx, axis=1, dtype=np.float32)  # noqa: E741
    return m, Z, A, I


def softargmax_combine(parts):
    """Merge a list of ``(m, Z, A, I)`` partials: returns global ``m, Z`` and ``depth = A/Z``, ``eidx = I/Z``."""
    m = np.max(np.stack([p[0] for p in parts]), axis=0)
    Z = np.zeros_like(m)
    A = np.zeros_like(m)
    I = np.zeros_like(m)  # noqa: E741
    for pm, pz, pa, pi in parts:
        s = np.exp(pm - m, dtype=np.float32)
        Z = Z + pz * s
        A = A + pa * s
        I = I + pi * s  # noqa: E741
    return m, Z, (A / Z).astype(np.float32), (I / Z).astype(np.float32)


def soft_argmin(volume, dim=1, keepdim=False, window=None):
    """utils.py:51-64."""
    p = _softmax(volume, dim)
    length = p.shape[dim]
    shape = [length if i == dim else 1 for i in range(p.ndim)]
    index = np.arange(length, dtype=np.float32).reshape(shape)
    out = np.sum(index * p, axis=dim, keepdims=True, dtype=np.float32)
    out_sq = out if keepdim else np.squeeze(out, dim)
    if window is None:
        return p, out_sq
    mask = (np.abs(index - out) <= window).astype(np.float32)
    prob_map = np.sum(p * mask, axis=dim, keepdims=keepdim, dtype=np.float32)
    return p, out_sq, prob_map


def entropy(volume, dim=1, keepdim=False):
    """utils.py:67-68."""
    v = _f32(volume)
    return np.sum(-v * np.log(np.clip(v, F32(1e-9), F32(1.0))), axis=dim, keepdims=keepdim, dtype=np.float32)


def softargmax_confidence_part(cost_slab, merged_m, merged_Z, merged_eidx, num_planes, d_begin):
    """One slab's share of the 4-bin confidence (mvsnet.py:143-158) around the GLOBAL index ``trunc(eidx)``:
    ``sum_{d in {idx-1..idx+2} & slab} exp(x_d - m) / Z``; the shares of all slabs sum to the confidence."""
    x = _f32(cost_slab)
    B, Dl, h, w = x.shape
    idx = np.clip(merged_eidx.astype(np.int64), 0, num_planes - 1)
    out = np.zeros((B, h, w), np.float32)
    for k in (-1, 0, 1, 2):
        d = idx + k
        inside = (d >= max(d_begin, 0)) & (d < min(d_begin + Dl, num_planes))
        local = np.clip(d - d_begin, 0, Dl - 1)
        val = np.exp(np.take_along_axis(x, local[:, None], axis=1)[:, 0] - merged_m, dtype=np.float32)
        out = out + np.where(inside, val, F32(0.0))
    return (out / merged_Z).astype(np.float32)
